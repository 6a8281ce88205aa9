#!/bin/bash
# The driver's command at N GPUs (default workload, other_configs included), with a readable digest.
# gpurun --gpus N --timeout 1500 -- bash tools/gpu_default_bench.sh N <tag>
N=${1:-1}; TAG=${2:-r03}
O=gpurun_out; mkdir -p $O
if [ "$N" -gt 1 ]; then
  time timeout 1400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus $N --steps 20 --warmup 5 > $O/${TAG}_bench_n$N.json 2> $O/${TAG}_bench_n$N.err
else
  time timeout 1400 python bench.py --steps 20 --warmup 5 > $O/${TAG}_bench_n$N.json 2> $O/${TAG}_bench_n$N.err
fi
echo "bench exit $?"
python - $O/${TAG}_bench_n$N.json <<'PY'
import json,sys
for l in open(sys.argv[1]):
    if l.startswith('{'):
        d=json.loads(l); print('headline', d['ms_per_step'], d['value'], d['clocks'], 'e2e', (d.get('e2e') or {}).get('value'))
        for k,v in (d.get('other_configs') or {}).items():
            c=(v.get('checks') or {}); s=c.get('joint_coverage_vs_cpu_arm') or {}
            print(' ', k, {x: v.get(x) for x in ('ms_per_step','value','wall_s','clocks','error','skipped')}, (v.get('config') or {}).get('global_batch'))
            print('     coverage', c.get('empirical_coverage'), 'torch-path', c.get('calibration_equals_independent_torch_path'), 'slab cpu/gpu', s.get('coverage_cpu_arm'), s.get('coverage_gpu'), s.get('coverage_equals_cpu_arm'), s.get('error'))
            print('     roofline', v.get('roofline'), 'e2e', v.get('e2e') if not isinstance(v.get('e2e'), dict) else v['e2e'].get('value', v['e2e']))
PY
tail -3 $O/${TAG}_bench_n$N.err | cut -c1-300
