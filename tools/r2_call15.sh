#!/bin/bash
mkdir -p gpurun_out/c15
timeout 600 python -m pytest tests/test_gpu_aten_cuda.py tests/test_gpu_local_corr.py -q -m gpu -s > gpurun_out/c15/pytest.log 2>&1
grep -E "aten-cuda|passed|failed|Error" gpurun_out/c15/pytest.log | cut -c1-250
