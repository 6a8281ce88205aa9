#!/bin/bash
# Round-2 second GPU pass (1 GPU): ncu launch lists of the Joint workloads + full captures of the MHD kernels.
TAG=${1:-r2b}
O=gpurun_out
mkdir -p $O
for w in mhd wave-joint burgers; do
  SHORT="python bench.py --workload $w --steps 2 --warmup 3 --no-e2e --no-cpu"
  timeout 300 $SHORT > $O/${TAG}_plain_$w.log 2>&1 &&
  timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off -c 400 --csv --log-file $O/${TAG}_launches_$w.csv $SHORT > $O/${TAG}_ncu_list_$w.log 2>&1
  echo "ncu list $w exit $?"
done
SHORT="python bench.py --workload mhd --steps 2 --warmup 3 --no-e2e --no-cpu"
timeout 900 ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:'star_kernel|cell_moments|joint_score' -s 12 -c 4 -o $O/${TAG}_prof_mhd $SHORT > $O/${TAG}_ncu_full_mhd.log 2>&1
echo "ncu full exit $?"
ncu -i $O/${TAG}_prof_mhd.ncu-rep --page raw --csv > $O/${TAG}_prof_mhd_raw.csv 2>/dev/null
timeout 600 python -m pytest tests -m gpu -x -q -k "autograd or grad" > $O/${TAG}_pytest_grad.log 2>&1; tail -5 $O/${TAG}_pytest_grad.log
ls -la $O | tail -12
