#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q -s > gpurun_out/c33_alone.log 2>&1; echo "rc=$?" >> gpurun_out/c33_alone.log
tail -n 6 gpurun_out/c33_alone.log
timeout 1800 python -m pytest tests -m gpu -x -q > gpurun_out/c33_all.log 2>&1; echo "rc=$?" >> gpurun_out/c33_all.log
tail -n 5 gpurun_out/c33_all.log
