#!/bin/bash
# Round-2 sixth GPU pass (1 GPU): the two tests that failed in pass r2e, with their tracebacks.
TAG=${1:-r2f}
O=gpurun_out
mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_boundary_spectral.py tests/test_gpu_stencil.py -m gpu -q --tb=short -k "periodic_boundary or flat_line or ragged" > $O/${TAG}_pytest.log 2>&1; echo "pytest exit $?" >> $O/${TAG}_pytest.log
tail -80 $O/${TAG}_pytest.log | cut -c1-400
