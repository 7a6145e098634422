#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/c25*
timeout 1200 python -m pytest tests/test_gpu_corr.py tests/test_gpu_aten_cuda.py tests/test_gpu_guard_words.py -m gpu -x -q > gpurun_out/c25_pytest.log 2>&1; echo "rc=$?" >> gpurun_out/c25_pytest.log
tail -n 4 gpurun_out/c25_pytest.log
timeout 600 python tools/kbench.py cfg2 --graph --only lookup > gpurun_out/c25_kb_cfg2.log 2>&1; grep -i "lookup" gpurun_out/c25_kb_cfg2.log
timeout 600 python tools/kbench.py cfg4 --graph --only lookup > gpurun_out/c25_kb_cfg4.log 2>&1; grep -i "lookup" gpurun_out/c25_kb_cfg4.log
