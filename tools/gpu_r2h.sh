#!/bin/bash
# Round-2 eighth GPU pass (1 GPU): flat-line kernel tests (ragged rows, JOREK stationary plane), generic-path bench.
TAG=${1:-r2h}
O=gpurun_out
mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_stencil.py -m gpu -q --tb=short -k "flat_line or jorek or ragged or decode" > $O/${TAG}_pytest.log 2>&1; echo "pytest exit $?" >> $O/${TAG}_pytest.log
tail -40 $O/${TAG}_pytest.log | cut -c1-400
timeout 300 python tools/generic_bench.py > $O/${TAG}_generic_bench.json 2> $O/${TAG}_generic_bench.err; echo "generic bench exit $?"; cat $O/${TAG}_generic_bench.json; tail -3 $O/${TAG}_generic_bench.err
