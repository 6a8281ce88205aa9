#!/bin/bash
# Round-2 third GPU pass (1 GPU): tests, A/B of the TMA L2 hints, benches of every workload.
TAG=${1:-r2c}
O=gpurun_out
mkdir -p $O
timeout 1200 python -m pytest tests -m gpu -x -q > $O/${TAG}_pytest_gpu.log 2>&1; echo "pytest exit $?" >> $O/${TAG}_pytest_gpu.log
tail -5 $O/${TAG}_pytest_gpu.log
for h in 0 1 2 3; do
  CPPRE_TMA_L2_HINTS=$h timeout 300 python bench.py --steps 10 --warmup 3 --no-e2e --no-cpu > $O/${TAG}_hints${h}_ns.json 2>> $O/${TAG}_hints.err
  CPPRE_TMA_L2_HINTS=$h timeout 300 python bench.py --workload mhd --steps 5 --warmup 3 --no-e2e --no-cpu > $O/${TAG}_hints${h}_mhd.json 2>> $O/${TAG}_hints.err
  CPPRE_TMA_L2_HINTS=$h timeout 300 python bench.py --workload wave-joint --steps 10 --warmup 3 --no-e2e --no-cpu > $O/${TAG}_hints${h}_wavej.json 2>> $O/${TAG}_hints.err
done
for w in burgers wave-joint; do
  timeout 600 python bench.py --workload $w --steps 20 --warmup 3 --no-e2e --no-cpu > $O/${TAG}_bench_$w.json 2> $O/${TAG}_bench_$w.err; echo "bench $w exit $?"
done
ls $O | grep ${TAG} | head -40
timeout 300 python tools/dropin_script_bench.py 32 > $O/${TAG}_dropin.json 2> $O/${TAG}_dropin.err; cat $O/${TAG}_dropin.json
