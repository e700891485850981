#!/bin/bash
# the no-allocation entry point: its test alone in a fresh process, then after the whole GPU suite (pytest -x order)
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q -x -p no:cacheprovider 2>&1 | tail -5
timeout 1800 python -m pytest tests -m gpu -q -x -p no:cacheprovider > gpurun_out/r2_pytest7.log 2>&1; tail -5 gpurun_out/r2_pytest7.log
