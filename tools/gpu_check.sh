#!/bin/bash
# One GPU-box pass: parity tests, the bench line, the ncu launch list and one full capture of
# the K1 / K3 kernels.  Run as: gpurun --timeout 1500 -- bash tools/gpu_check.sh <tag>
TAG=${1:-run}
O=gpurun_out
mkdir -p $O
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $O/${TAG}_smi.txt 2>&1
timeout 900 python -m pytest tests -m gpu -x -q > $O/${TAG}_pytest_gpu.log 2>&1; echo "pytest exit $?" >> $O/${TAG}_pytest_gpu.log
tail -3 $O/${TAG}_pytest_gpu.log
timeout 600 python bench.py --steps 10 --warmup 3 > $O/${TAG}_bench.json 2> $O/${TAG}_bench.err; echo "bench exit $?"
cat $O/${TAG}_bench.json
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > $O/${TAG}_bench_ref.json 2> $O/${TAG}_bench_ref.err; echo "ref exit $?"
cat $O/${TAG}_bench_ref.json
SHORT="python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu"
timeout 300 $SHORT > $O/${TAG}_plain.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off -c 400 --csv --log-file $O/${TAG}_launches.csv $SHORT > $O/${TAG}_ncu_list.log 2>&1
echo "ncu list exit $?"
timeout 300 $SHORT > $O/${TAG}_plain2.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:'star_kernel|select_' -s 9 -c 3 -o $O/${TAG}_prof $SHORT > $O/${TAG}_ncu_full.log 2>&1
echo "ncu full exit $?"
ls -la $O
