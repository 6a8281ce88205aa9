#!/bin/bash
# Round-2 fourth GPU pass (1 GPU): the t-innermost (FNO layout) star kernel -- parity tests, memcheck of a small
# case, throughput per layout, launch list.
TAG=${1:-r2d}
O=gpurun_out
mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_stencil.py -m gpu -x -q -k "innermost or fno" > $O/${TAG}_pytest_tin.log 2>&1; echo "pytest exit $?" >> $O/${TAG}_pytest_tin.log
tail -15 $O/${TAG}_pytest_tin.log
timeout 600 compute-sanitizer --tool memcheck --error-exitcode 7 python tools/tin_small.py > $O/${TAG}_memcheck.log 2>&1; echo "memcheck exit $?" | tee -a $O/${TAG}_memcheck.log
tail -4 $O/${TAG}_memcheck.log
timeout 600 python tools/perm_layout.py > $O/${TAG}_perm_layout.json 2> $O/${TAG}_perm_layout.err; echo "perm_layout exit $?"; cat $O/${TAG}_perm_layout.json; tail -3 $O/${TAG}_perm_layout.err
timeout 900 python -m pytest tests -m gpu -x -q > $O/${TAG}_pytest_gpu.log 2>&1; echo "pytest exit $?" >> $O/${TAG}_pytest_gpu.log
tail -5 $O/${TAG}_pytest_gpu.log
# full captures: burgers STORE + COVER (F=1, batch-rows variant), and the t-innermost kernel
timeout 600 ncu --set full --clock-control none --import-source on -k regex:star_kernel --launch-skip 6 --launch-count 2 -f -o $O/${TAG}_prof_burgers python bench.py --workload burgers --steps 2 --warmup 3 --no-e2e --no-cpu > $O/${TAG}_prof_burgers.log 2>&1; echo "ncu burgers exit $?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:tin_kernel --launch-skip 2 --launch-count 3 -f -o $O/${TAG}_prof_tin python tools/perm_layout.py > $O/${TAG}_prof_tin.log 2>&1; echo "ncu tin exit $?"
ls -la $O | grep ${TAG}
