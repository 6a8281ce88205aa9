#!/bin/bash
# Final 1-GPU verification of the tree as it stands: whole GPU suite, smoke, default bench line (both arms).
TAG=${1:-final}
O=gpurun_out
mkdir -p $O
timeout 1200 python -m pytest tests -m gpu -q --tb=short > $O/${TAG}_pytest_gpu.log 2>&1; echo "pytest exit $?" >> $O/${TAG}_pytest_gpu.log
tail -6 $O/${TAG}_pytest_gpu.log | cut -c1-300
timeout 300 python __graft_entry__.py smoke > $O/${TAG}_smoke.log 2>&1; echo "smoke exit $?"; tail -1 $O/${TAG}_smoke.log
timeout 600 python bench.py > $O/${TAG}_bench_ns.json 2> $O/${TAG}_bench_ns.err; echo "bench exit $?"; cut -c1-300 $O/${TAG}_bench_ns.json; tail -2 $O/${TAG}_bench_ns.err
timeout 900 python bench.py --impl reference --steps 2 --warmup 1 > $O/${TAG}_bench_ref.json 2> $O/${TAG}_bench_ref.err; echo "ref exit $?"; cut -c1-300 $O/${TAG}_bench_ref.json
timeout 300 python bench.py --workload wave --steps 10 --warmup 3 --no-e2e > $O/${TAG}_bench_wave.json 2> $O/${TAG}_bench_wave.err; echo "wave exit $?"; cut -c1-200 $O/${TAG}_bench_wave.json; tail -2 $O/${TAG}_bench_wave.err
