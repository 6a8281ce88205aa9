#!/bin/bash
# Round-2 first GPU pass (1 GPU): parity tests, NS bench, MHD / wave-joint / burgers lines, reference arm.
TAG=${1:-r2a}
O=gpurun_out
mkdir -p $O
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $O/${TAG}_smi.txt 2>&1
nproc > $O/${TAG}_host.txt; free -g >> $O/${TAG}_host.txt; nvidia-smi topo -m >> $O/${TAG}_host.txt 2>&1
timeout 1200 python -m pytest tests -m gpu -x -q > $O/${TAG}_pytest_gpu.log 2>&1; echo "pytest exit $?" >> $O/${TAG}_pytest_gpu.log
tail -5 $O/${TAG}_pytest_gpu.log
timeout 600 python bench.py --steps 10 --warmup 3 > $O/${TAG}_bench_ns.json 2> $O/${TAG}_bench_ns.err; echo "bench ns exit $?"
cat $O/${TAG}_bench_ns.json
for eq in momentum energy induction continuity gauss; do
  timeout 600 python bench.py --workload mhd --equation $eq --steps 5 --warmup 3 --no-e2e --no-cpu > $O/${TAG}_bench_mhd_$eq.json 2> $O/${TAG}_bench_mhd_$eq.err; echo "bench mhd $eq exit $?"
  cat $O/${TAG}_bench_mhd_$eq.json
done
for w in wave-joint burgers; do
  timeout 600 python bench.py --workload $w --steps 10 --warmup 3 > $O/${TAG}_bench_$w.json 2> $O/${TAG}_bench_$w.err; echo "bench $w exit $?"
  cat $O/${TAG}_bench_$w.json
done
timeout 900 python bench.py --impl reference --steps 2 --warmup 1 --no-full-cpu > $O/${TAG}_bench_ref.json 2> $O/${TAG}_bench_ref.err; echo "ref exit $?"
cat $O/${TAG}_bench_ref.json
ls -la $O | tail -30
