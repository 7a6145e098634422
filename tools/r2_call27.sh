#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/c27*
timeout 1200 python -m pytest tests/test_gpu_corr.py tests/test_gpu_aten_cuda.py tests/test_gpu_guard_words.py tests/test_gpu_autograd_r2.py -m gpu -x -q > gpurun_out/c27_pytest.log 2>&1; echo "rc=$?" >> gpurun_out/c27_pytest.log
tail -n 4 gpurun_out/c27_pytest.log
timeout 600 python tools/kbench.py cfg2 --graph --only lookup > gpurun_out/c27_kb_cfg2.log 2>&1; grep -i "lookup" gpurun_out/c27_kb_cfg2.log
timeout 600 python tools/kbench.py cfg4 --graph --only lookup > gpurun_out/c27_kb_cfg4.log 2>&1; grep -i "lookup" gpurun_out/c27_kb_cfg4.log
timeout 900 ncu --set full --clock-control none --import-source on -k regex:lookup_kernel -s 20 -c 1 -o gpurun_out/c27_lookup python tools/kbench.py cfg2 --only lookup --reps 3 > gpurun_out/c27_ncu.log 2>&1; tail -n 2 gpurun_out/c27_ncu.log
python tools/ncu_brief.py gpurun_out/c27_lookup.ncu-rep > gpurun_out/c27_ncu_brief.txt 2>&1; head -40 gpurun_out/c27_ncu_brief.txt
