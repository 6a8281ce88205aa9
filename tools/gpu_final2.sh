#!/bin/bash
# Last 1-GPU check of the committed tree: whole GPU suite, smoke, a short bench line.
TAG=${1:-final2}
O=gpurun_out
mkdir -p $O
timeout 1200 python -m pytest tests -m gpu -q --tb=short > $O/${TAG}_pytest_gpu.log 2>&1; echo "pytest exit $?" >> $O/${TAG}_pytest_gpu.log
tail -4 $O/${TAG}_pytest_gpu.log | cut -c1-300
timeout 300 python __graft_entry__.py smoke > $O/${TAG}_smoke.log 2>&1; echo "smoke exit $?"; tail -1 $O/${TAG}_smoke.log
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu > $O/${TAG}_bench_ns.json 2> $O/${TAG}_bench_ns.err; echo "bench exit $?"; cut -c1-260 $O/${TAG}_bench_ns.json
timeout 600 python tools/perm_layout.py > $O/${TAG}_perm_layout.json 2>/dev/null; cut -c1-300 $O/${TAG}_perm_layout.json
timeout 300 python tools/generic_bench.py > $O/${TAG}_generic_bench.json 2>/dev/null; python -c "
import json; d=json.load(open('$O/${TAG}_generic_bench.json'))
print({k:(v['path'],v['gpoints_s']) for k,v in d.items()})"
