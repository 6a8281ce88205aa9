#!/bin/bash
# ncu capture of the t-innermost star kernel (K1t) and the flat-line kernel (K1f) on their throughput tools.
TAG=${1:-tinprof}
O=gpurun_out
mkdir -p $O
timeout 600 ncu --set full --clock-control none --import-source on -k regex:tin_kernel --launch-skip 2 --launch-count 2 -f -o /tmp/${TAG}_tin python tools/perm_layout.py > $O/${TAG}_tin.log 2>&1; echo "ncu tin exit $?"
ncu -i /tmp/${TAG}_tin.ncu-rep --page raw --csv > $O/${TAG}_tin_raw.csv 2>/dev/null
python tools/ncu_summary.py $O/${TAG}_tin_raw.csv > $O/${TAG}_tin_summary.txt 2>&1
ncu -i /tmp/${TAG}_tin.ncu-rep --page details --csv 2>/dev/null | grep -i -E "stall|Warp Cycles Per|No Eligible|Issued Warp|Bank conflict|Registers|Theoretical Occ|Achieved Occ|L1/TEX Hit|L2 Hit|Mem Busy|Max Bandwidth" | cut -c1-260 | head -60 > $O/${TAG}_tin_details.txt
cat $O/${TAG}_tin_summary.txt | head -40
ncu -i /tmp/${TAG}_tin.ncu-rep --page source --csv 2>/dev/null | head -3 | cut -c1-600
ncu -i /tmp/${TAG}_tin.ncu-rep --page raw --csv 2>/dev/null | python -c "
import csv,sys
rows=list(csv.reader(sys.stdin)); h=rows[0]
want=[i for i,c in enumerate(h) if 'smsp__pcsamp_warps_issue_stalled' in c or 'smsp__average_warp' in c or 'shared_ld' in c]
for r in rows[2:3]:
    for i in want: print(h[i][:90], r[i])
" > $O/${TAG}_tin_stalls.txt 2>&1
cat $O/${TAG}_tin_stalls.txt | sort -k2 -n -r | head -30
