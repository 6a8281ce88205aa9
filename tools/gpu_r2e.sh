#!/bin/bash
# Round-2 fifth GPU pass (1 GPU): whole GPU suite, smoke, the default bench line + reference arm, ncu launch
# list and full captures of the headline kernels and of the MHD star kernel, FNO-layout throughput.
# gpurun --timeout 1500 -- bash tools/gpu_r2e.sh <tag>
TAG=${1:-r2e}
O=gpurun_out
mkdir -p $O
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $O/${TAG}_smi.txt 2>&1
timeout 1200 python -m pytest tests -m gpu -q > $O/${TAG}_pytest_gpu.log 2>&1; echo "pytest exit $?" >> $O/${TAG}_pytest_gpu.log
tail -4 $O/${TAG}_pytest_gpu.log
timeout 300 python __graft_entry__.py smoke > $O/${TAG}_smoke.log 2>&1; echo "smoke exit $?"; tail -1 $O/${TAG}_smoke.log
timeout 600 python bench.py --steps 10 --warmup 3 > $O/${TAG}_bench.json 2> $O/${TAG}_bench.err; echo "bench exit $?"
cut -c1-1200 $O/${TAG}_bench.json
SHORT="python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu"
timeout 300 $SHORT > $O/${TAG}_plain.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off -c 400 --csv --log-file $O/${TAG}_launches.csv $SHORT > $O/${TAG}_ncu_list.log 2>&1
echo "ncu list exit $?"
timeout 900 ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:'star_kernel|select_' -s 9 -c 3 -f -o $O/${TAG}_prof $SHORT > $O/${TAG}_ncu_full.log 2>&1
echo "ncu full exit $?"
ncu -i $O/${TAG}_prof.ncu-rep --page raw --csv > $O/${TAG}_prof_raw.csv 2>/dev/null
python tools/ncu_summary.py $O/${TAG}_prof_raw.csv > $O/${TAG}_prof_summary.txt 2>&1
timeout 600 python tools/perm_layout.py > $O/${TAG}_perm_layout.json 2> $O/${TAG}_perm_layout.err; echo "perm_layout exit $?"; cat $O/${TAG}_perm_layout.json; tail -3 $O/${TAG}_perm_layout.err
MHD="python bench.py --workload mhd --global-batch 128 --steps 2 --warmup 3 --no-e2e --no-cpu"
timeout 300 python bench.py --workload mhd --global-batch 256 --steps 10 --warmup 3 > $O/${TAG}_bench_mhd.json 2> $O/${TAG}_bench_mhd.err; echo "bench mhd exit $?"
cut -c1-600 $O/${TAG}_bench_mhd.json
timeout 300 python bench.py --workload advection --steps 20 --warmup 3 > $O/${TAG}_bench_advection.json 2> $O/${TAG}_bench_advection.err; echo "bench advection exit $?"; cut -c1-700 $O/${TAG}_bench_advection.json; tail -3 $O/${TAG}_bench_advection.err
timeout 900 ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:'star_kernel|cell_|joint_score' -s 10 -c 5 -f -o $O/${TAG}_prof_mhd $MHD > $O/${TAG}_ncu_full_mhd.log 2>&1
echo "ncu mhd exit $?"
ncu -i $O/${TAG}_prof_mhd.ncu-rep --page raw --csv > $O/${TAG}_prof_mhd_raw.csv 2>/dev/null
python tools/ncu_summary.py $O/${TAG}_prof_mhd_raw.csv > $O/${TAG}_prof_mhd_summary.txt 2>&1
timeout 300 python tools/generic_bench.py > $O/${TAG}_generic_bench.json 2> $O/${TAG}_generic_bench.err; echo "generic bench exit $?"; cat $O/${TAG}_generic_bench.json | cut -c1-2500; tail -3 $O/${TAG}_generic_bench.err
timeout 300 python tools/k3_sweep.py 256,512,2048,4096 > $O/${TAG}_k3_sweep.jsonl 2> $O/${TAG}_k3_sweep.err; echo "k3 sweep exit $?"; cat $O/${TAG}_k3_sweep.jsonl
for w in burgers wave-joint; do
  timeout 600 python bench.py --workload $w --steps 20 --warmup 3 > $O/${TAG}_bench_$w.json 2> $O/${TAG}_bench_$w.err; echo "bench $w exit $?"
done
timeout 600 python bench.py --impl reference --steps 1 --warmup 0 > $O/${TAG}_bench_ref.json 2> $O/${TAG}_bench_ref.err; echo "ref exit $?"
cut -c1-1500 $O/${TAG}_bench_ref.json
ls -la $O | grep ${TAG}
