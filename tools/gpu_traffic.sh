#!/bin/bash
# Re-captures the DRAM traffic of the K1 STORE / COVER launches (ncu --set full) and the launch list of the NS step, so that
# bench.py's roofline.traffic belongs to the kernel sources of THIS tree.  gpurun --timeout 1800 -- bash tools/gpu_traffic.sh <tag>
TAG=${1:-r03}
O=gpurun_out; mkdir -p $O
SHORT="python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu"
timeout 300 $SHORT > $O/${TAG}_short.json 2> $O/${TAG}_short.err; echo "short bench exit $?"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off -c 400 --csv --log-file $O/${TAG}_launches.csv $SHORT > $O/${TAG}_ncu_list.log 2>&1; echo "ncu list exit $?"
timeout 900 ncu --set full --clock-control none --profile-from-start off -k regex:'star_kernel|select_' -s 9 -c 3 -f -o /tmp/${TAG}_prof $SHORT > $O/${TAG}_ncu_full.log 2>&1; echo "ncu full exit $?"
ncu -i /tmp/${TAG}_prof.ncu-rep --page raw --csv > $O/${TAG}_prof_raw.csv 2>/dev/null
python tools/ncu_summary.py $O/${TAG}_prof_raw.csv > $O/${TAG}_prof_summary.txt 2>&1
python tools/k1_traffic.py $O/${TAG}_prof_raw.csv ns "" 256 && cp profiles/k1_traffic_ns.json $O/${TAG}_k1_traffic_ns.json
