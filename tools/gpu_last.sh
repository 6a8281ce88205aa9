#!/bin/bash
# Last check (2 GPUs): whole GPU suite incl. the two-device test, smoke, NS at 2 GPUs, both arms at 1 GPU (short).
O=gpurun_out; mkdir -p $O
timeout 1200 python -m pytest tests -m gpu -q --tb=short > $O/last_pytest_gpu.log 2>&1; echo "pytest exit $?" >> $O/last_pytest_gpu.log; tail -3 $O/last_pytest_gpu.log | cut -c1-200
timeout 300 python __graft_entry__.py smoke 2>&1 | tail -1
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 --no-cpu --no-e2e > $O/last_bench_ns_n2.json 2> $O/last_bench_ns_n2.err; echo "ns n2 exit $?"; cut -c1-200 $O/last_bench_ns_n2.json
