#!/bin/bash
# 8-GPU pass of the final tree: NS weak-scaling points (8, 4), config 5 at full size, K3x phase profile.
# gpurun --gpus 8 --timeout 900 -- bash tools/gpu_n8b.sh <tag>
TAG=${1:-n8b}
O=gpurun_out
mkdir -p $O
TR() { python -m torch.distributed.run --nnodes=1 --nproc-per-node $1 --master-addr 127.0.0.1 --master-port 29511 "${@:2}"; }
timeout 400 bash -c "$(declare -f TR); TR 8 bench.py --gpus 8 --steps 10 --warmup 3 --no-cpu" > $O/${TAG}_ns_n8.json 2> $O/${TAG}_ns_n8.err; echo "ns n8 exit $?"; tail -2 $O/${TAG}_ns_n8.err | cut -c1-300
timeout 400 bash -c "$(declare -f TR); TR 8 bench.py --gpus 8 --workload mhd --steps 10 --warmup 3 --no-cpu --no-e2e" > $O/${TAG}_mhd_full_n8.json 2> $O/${TAG}_mhd_full_n8.err; echo "mhd full n8 exit $?"; tail -2 $O/${TAG}_mhd_full_n8.err | cut -c1-300
timeout 300 bash -c "$(declare -f TR); TR 4 bench.py --gpus 4 --steps 10 --warmup 3 --no-cpu --no-e2e" > $O/${TAG}_ns_n4.json 2> $O/${TAG}_ns_n4.err; echo "ns n4 exit $?"
timeout 200 bash -c "$(declare -f TR); TR 8 tools/dist_profile.py" > $O/${TAG}_dist_profile_n8.json 2> $O/${TAG}_dist_profile_n8.err; echo "profile n8 exit $?"
cat $O/${TAG}_*.json | cut -c1-330
