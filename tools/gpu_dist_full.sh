#!/bin/bash
# gpurun --gpus N -- bash tools/gpu_dist_full.sh <tag> <N>: GPU parity tests on one GPU, then the multi-GPU pass
TAG=${1:-d}; N=${2:-2}; O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q > $O/${TAG}_pytest_gpu.log 2>&1; echo "pytest exit $?" >> $O/${TAG}_pytest_gpu.log
tail -4 $O/${TAG}_pytest_gpu.log
bash tools/gpu_dist.sh $TAG $N
