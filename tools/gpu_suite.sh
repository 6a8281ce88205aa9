#!/bin/bash
# Whole GPU suite + smoke on one GPU.
O=gpurun_out; mkdir -p $O
# the no-alloc entry point must also pass alone in a fresh process (it does its own per-device set-up)
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q --tb=short 2>&1 | tail -1
timeout 1200 python -m pytest tests -m gpu -q --tb=short > $O/suite_pytest_gpu.log 2>&1; echo "pytest exit $?" >> $O/suite_pytest_gpu.log; tail -3 $O/suite_pytest_gpu.log | cut -c1-200
timeout 300 python __graft_entry__.py smoke 2>&1 | tail -1
