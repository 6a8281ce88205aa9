#!/bin/bash
# Whole GPU suite + smoke on one GPU.
O=gpurun_out; mkdir -p $O
timeout 1200 python -m pytest tests -m gpu -q --tb=short > $O/suite_pytest_gpu.log 2>&1; echo "pytest exit $?" >> $O/suite_pytest_gpu.log; tail -3 $O/suite_pytest_gpu.log | cut -c1-200
timeout 300 python __graft_entry__.py smoke 2>&1 | tail -1
