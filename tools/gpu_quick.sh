#!/bin/bash
# Quick GPU pass: parity tests, short bench, kernel sweeps.  gpurun --timeout 900 -- bash tools/gpu_quick.sh <tag>
TAG=${1:-q}
O=gpurun_out
mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q > $O/${TAG}_pytest_gpu.log 2>&1; echo "pytest exit $?" >> $O/${TAG}_pytest_gpu.log
tail -5 $O/${TAG}_pytest_gpu.log
timeout 600 python bench.py --steps 10 --warmup 3 --no-e2e --no-cpu > $O/${TAG}_bench.json 2> $O/${TAG}_bench.err; echo "bench exit $?"
cat $O/${TAG}_bench.json
timeout 300 python tools/k3_sweep.py > $O/${TAG}_k3_sweep.jsonl 2>&1; cat $O/${TAG}_k3_sweep.jsonl
if [ -n "$K1SWEEP" ]; then timeout 900 python tools/k1_sweep.py $K1SWEEP > $O/${TAG}_k1_sweep.jsonl 2>&1; cat $O/${TAG}_k1_sweep.jsonl; fi
