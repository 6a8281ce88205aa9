#!/bin/bash
# Round-2 seventh GPU pass (1 GPU): whole GPU suite, smoke, every workload's bench line, ncu launch list and full
# captures (raw CSV kept, .ncu-rep dropped: gpurun_out/ must stay under 64 MiB), layout / generic-path throughput.
TAG=${1:-r2g}
O=gpurun_out
mkdir -p $O
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $O/${TAG}_smi.txt 2>&1
timeout 1200 python -m pytest tests -m gpu -q --tb=short > $O/${TAG}_pytest_gpu.log 2>&1; echo "pytest exit $?" >> $O/${TAG}_pytest_gpu.log
tail -30 $O/${TAG}_pytest_gpu.log | cut -c1-300
timeout 300 python __graft_entry__.py smoke > $O/${TAG}_smoke.log 2>&1; echo "smoke exit $?"; tail -1 $O/${TAG}_smoke.log
timeout 600 python bench.py --steps 10 --warmup 3 > $O/${TAG}_bench_ns.json 2> $O/${TAG}_bench_ns.err; echo "bench exit $?"
SHORT="python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off -c 400 --csv --log-file $O/${TAG}_launches.csv $SHORT > $O/${TAG}_ncu_list.log 2>&1
echo "ncu list exit $?"
timeout 900 ncu --set full --clock-control none --profile-from-start off -k regex:'star_kernel|select_' -s 9 -c 3 -f -o /tmp/${TAG}_prof $SHORT > $O/${TAG}_ncu_full.log 2>&1
echo "ncu full exit $?"
ncu -i /tmp/${TAG}_prof.ncu-rep --page raw --csv > $O/${TAG}_prof_raw.csv 2>/dev/null
python tools/ncu_summary.py $O/${TAG}_prof_raw.csv > $O/${TAG}_prof_summary.txt 2>&1
MHD="python bench.py --workload mhd --global-batch 128 --steps 2 --warmup 3 --no-e2e --no-cpu"
timeout 900 ncu --set full --clock-control none --profile-from-start off -k regex:'star_kernel|cell_|joint_score' -s 10 -c 5 -f -o /tmp/${TAG}_prof_mhd $MHD > $O/${TAG}_ncu_full_mhd.log 2>&1
echo "ncu mhd exit $?"
ncu -i /tmp/${TAG}_prof_mhd.ncu-rep --page raw --csv > $O/${TAG}_prof_mhd_raw.csv 2>/dev/null
python tools/ncu_summary.py $O/${TAG}_prof_mhd_raw.csv > $O/${TAG}_prof_mhd_summary.txt 2>&1
timeout 300 python bench.py --workload mhd --global-batch 256 --steps 10 --warmup 3 > $O/${TAG}_bench_mhd.json 2> $O/${TAG}_bench_mhd.err; echo "bench mhd exit $?"
timeout 300 python bench.py --workload advection --steps 20 --warmup 3 > $O/${TAG}_bench_advection.json 2> $O/${TAG}_bench_advection.err; echo "bench advection exit $?"
for w in burgers wave-joint; do
  timeout 600 python bench.py --workload $w --steps 20 --warmup 3 > $O/${TAG}_bench_$w.json 2> $O/${TAG}_bench_$w.err; echo "bench $w exit $?"
done
timeout 600 python tools/perm_layout.py > $O/${TAG}_perm_layout.json 2> $O/${TAG}_perm_layout.err; echo "perm_layout exit $?"
timeout 300 python tools/generic_bench.py > $O/${TAG}_generic_bench.json 2> $O/${TAG}_generic_bench.err; echo "generic bench exit $?"; cut -c1-1500 $O/${TAG}_generic_bench.json
for f in ns mhd advection burgers wave-joint; do echo "== $f"; cut -c1-330 $O/${TAG}_bench_$f.json; done
du -sh $O; ls -la $O | grep ${TAG}
