#!/bin/bash
# K3x pass on ONE GPU: parity tests, short bench, the fused-exchange kernels with the ranks emulated (tools/k3x_profile.py), ncu of window / owner.
TAG=${1:-k3x}; O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q > $O/${TAG}_pytest_gpu.log 2>&1; echo "pytest exit $?" >> $O/${TAG}_pytest_gpu.log
tail -4 $O/${TAG}_pytest_gpu.log
timeout 600 python bench.py --steps 10 --warmup 3 --no-e2e --no-cpu > $O/${TAG}_bench.json 2> $O/${TAG}_bench.err; echo "bench exit $?"
python -c "
import json; d=json.load(open('$O/${TAG}_bench.json')); print(d['ms_per_step'], d['roofline']['kernel_ms'], d['roofline']['cover_kernel']['kernel_ms'])"
for w in 2 8; do timeout 300 python tools/k3x_profile.py $w > $O/${TAG}_k3x_w$w.json 2>&1; cat $O/${TAG}_k3x_w$w.json; done
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'window_exchange' -s 2 -c 1 -o $O/${TAG}_k3x_window python tools/k3x_profile.py 8 > $O/${TAG}_ncu_k3x.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'owner_select' -s 1 -c 1 -o $O/${TAG}_k3x_owner8 python tools/k3x_profile.py 8 >> $O/${TAG}_ncu_k3x.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'owner_select' -s 1 -c 1 -o $O/${TAG}_k3x_owner2 python tools/k3x_profile.py 2 >> $O/${TAG}_ncu_k3x.log 2>&1
echo "ncu k3x exit $?"
