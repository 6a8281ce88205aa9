#!/bin/bash
# 2-GPU validation of the final tree: gpurun --gpus 2 --timeout 900 -- bash tools/gpu_n2b.sh <tag>
TAG=${1:-n2b}
N=2
O=gpurun_out
mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
timeout 300 python -m pytest tests -m gpu -x -q -k "two_devices" > $O/${TAG}_pytest_2dev.log 2>&1; tail -2 $O/${TAG}_pytest_2dev.log
timeout 600 $TR bench.py --gpus $N --steps 10 --warmup 3 --no-cpu > $O/${TAG}_bench_ns.json 2> $O/${TAG}_bench_ns.err; echo "ns exit $?"; tail -3 $O/${TAG}_bench_ns.err
timeout 600 $TR bench.py --gpus $N --workload mhd --steps 5 --warmup 3 --no-cpu --no-e2e > $O/${TAG}_bench_mhd.json 2> $O/${TAG}_bench_mhd.err; echo "mhd exit $?"; tail -3 $O/${TAG}_bench_mhd.err
timeout 600 $TR bench.py --gpus $N --workload wave-joint --steps 10 --warmup 3 --no-cpu --no-e2e > $O/${TAG}_bench_wave-joint.json 2> $O/${TAG}_bench_wave-joint.err; echo "wave-joint exit $?"
timeout 300 $TR bench.py --impl reference --gpus $N --steps 1 --warmup 0 --no-full-cpu > $O/${TAG}_bench_ref.json 2> $O/${TAG}_bench_ref.err; echo "ref exit $?"
cat $O/${TAG}_bench_*.json | cut -c1-500
