#!/bin/bash
# 2-GPU validation pass: gpurun --gpus 2 --timeout 1500 -- bash tools/gpu_n2.sh <tag>
TAG=${1:-n2}
N=2
O=gpurun_out
mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
timeout 300 python -m pytest tests -m gpu -x -q -k "two_devices" > $O/${TAG}_pytest_2dev.log 2>&1; tail -2 $O/${TAG}_pytest_2dev.log
timeout 600 $TR bench.py --gpus $N --steps 10 --warmup 3 --no-cpu > $O/${TAG}_bench_ns.json 2> $O/${TAG}_bench_ns.err; echo "ns exit $?"; tail -3 $O/${TAG}_bench_ns.err
timeout 600 $TR bench.py --gpus $N --workload mhd --steps 5 --warmup 3 --no-cpu --no-e2e > $O/${TAG}_bench_mhd.json 2> $O/${TAG}_bench_mhd.err; echo "mhd exit $?"; tail -3 $O/${TAG}_bench_mhd.err
timeout 600 $TR bench.py --gpus $N --workload mhd --global-batch 128 --steps 5 --warmup 3 --no-cpu > $O/${TAG}_bench_mhd_g128.json 2> $O/${TAG}_bench_mhd_g128.err; echo "mhd g128 exit $?"; tail -3 $O/${TAG}_bench_mhd_g128.err
for w in wave-joint burgers; do
  timeout 600 $TR bench.py --gpus $N --workload $w --steps 10 --warmup 3 --no-cpu --no-e2e > $O/${TAG}_bench_$w.json 2> $O/${TAG}_bench_$w.err; echo "$w exit $?"; tail -3 $O/${TAG}_bench_$w.err
done
timeout 300 $TR tools/dist_profile.py > $O/${TAG}_dist_profile.json 2> $O/${TAG}_dist_profile.err; echo "profile exit $?"; cat $O/${TAG}_dist_profile.json
timeout 300 $TR tools/h2d_ceiling.py > $O/${TAG}_h2d.json 2> $O/${TAG}_h2d.err; cat $O/${TAG}_h2d.json
timeout 300 python tools/dropin_script_bench.py 32 > $O/${TAG}_dropin.json 2> $O/${TAG}_dropin.err; cat $O/${TAG}_dropin.json
cat $O/${TAG}_bench_*.json | cut -c1-600
