#!/bin/bash
# 8-GPU pass: config 5 (MHD Joint CP, BS=2048 over 8 GPUs), its strong-scaling record, the NS weak-scaling
# point + K3x phase profile, and the host's H2D ceiling.  gpurun --gpus 8 --timeout 1500 -- bash tools/gpu_n8.sh <tag>
TAG=${1:-n8}
O=gpurun_out
mkdir -p $O
TR() { python -m torch.distributed.run --nnodes=1 --nproc-per-node $1 --master-addr 127.0.0.1 --master-port 29511 "${@:2}"; }
nvidia-smi topo -m > $O/${TAG}_topo.txt 2>&1; nproc >> $O/${TAG}_topo.txt; free -g >> $O/${TAG}_topo.txt
timeout 600 bash -c "$(declare -f TR); TR 8 bench.py --gpus 8 --workload mhd --steps 10 --warmup 3 --no-cpu" > $O/${TAG}_mhd_full_n8.json 2> $O/${TAG}_mhd_full_n8.err; echo "mhd full n8 exit $?"; tail -2 $O/${TAG}_mhd_full_n8.err | cut -c1-300
for n in 1 2 4 8; do
  timeout 300 bash -c "$(declare -f TR); TR $n bench.py --gpus $n --workload mhd --global-batch 512 --steps 10 --warmup 3 --no-cpu --no-e2e" > $O/${TAG}_mhd_strong_n$n.json 2> $O/${TAG}_mhd_strong_n$n.err; echo "mhd strong n$n exit $?"
done
timeout 400 bash -c "$(declare -f TR); TR 4 bench.py --gpus 4 --workload mhd --steps 10 --warmup 3 --no-cpu --no-e2e" > $O/${TAG}_mhd_full_n4.json 2> $O/${TAG}_mhd_full_n4.err; echo "mhd full n4 exit $?"
timeout 400 bash -c "$(declare -f TR); TR 8 bench.py --gpus 8 --workload mhd --equation energy --steps 10 --warmup 3 --no-cpu --no-e2e" > $O/${TAG}_mhd_energy_n8.json 2> $O/${TAG}_mhd_energy_n8.err; echo "mhd energy n8 exit $?"
timeout 600 bash -c "$(declare -f TR); TR 8 bench.py --gpus 8 --steps 10 --warmup 3 --no-cpu" > $O/${TAG}_ns_n8.json 2> $O/${TAG}_ns_n8.err; echo "ns n8 exit $?"; tail -2 $O/${TAG}_ns_n8.err | cut -c1-300
timeout 300 bash -c "$(declare -f TR); TR 4 bench.py --gpus 4 --steps 10 --warmup 3 --no-cpu --no-e2e" > $O/${TAG}_ns_n4.json 2> $O/${TAG}_ns_n4.err; echo "ns n4 exit $?"
for n in 4 8; do
  timeout 200 bash -c "$(declare -f TR); TR $n tools/dist_profile.py" > $O/${TAG}_dist_profile_n$n.json 2> $O/${TAG}_dist_profile_n$n.err; echo "profile n$n exit $?"
  timeout 200 bash -c "$(declare -f TR); TR $n tools/h2d_ceiling.py" > $O/${TAG}_h2d_n$n.json 2> $O/${TAG}_h2d_n$n.err
done
CPPRE_KEYS_ALLREDUCE=0 timeout 200 bash -c "$(declare -f TR); TR 8 tools/dist_profile.py" > $O/${TAG}_dist_profile_n8_ka0.json 2>> $O/${TAG}_dist_profile_n8.err
timeout 300 bash -c "$(declare -f TR); TR 8 bench.py --gpus 8 --workload wave-joint --steps 10 --warmup 3 --no-cpu --no-e2e" > $O/${TAG}_wavej_n8.json 2> $O/${TAG}_wavej_n8.err; echo "wavej n8 exit $?"
cat $O/${TAG}_*.json | cut -c1-330
