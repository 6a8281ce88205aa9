#!/bin/bash
# Multi-GPU pass: per-phase profile of the marginal quantile exchange + the bench line at N GPUs.
# gpurun --gpus N --timeout 900 -- bash tools/gpu_dist.sh <tag> <N>
TAG=${1:-d}
N=${2:-2}
O=gpurun_out
mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
for KA in 0 1; do
CPPRE_KEYS_ALLREDUCE=$KA timeout 300 $TR tools/dist_profile.py > $O/${TAG}_dist_profile_n${N}_ka$KA.json 2> $O/${TAG}_dist_profile_n$N.err; echo "profile exit $?"
cat $O/${TAG}_dist_profile_n${N}_ka$KA.json
done
timeout 600 $TR bench.py --gpus $N --steps 10 --warmup 3 > $O/${TAG}_bench_n$N.json 2> $O/${TAG}_bench_n$N.err; echo "bench exit $?"
cat $O/${TAG}_bench_n$N.json
tail -5 $O/${TAG}_bench_n$N.err
