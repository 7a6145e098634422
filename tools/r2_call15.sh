#!/bin/bash
# staged warp A/B
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_warp.py tests/test_gpu_aten_cuda.py -m gpu -x -q > gpurun_out/c15_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/c15_pytest.log
for st in 0 1; do
  CF_WARP_STAGED=$st timeout 300 python tools/kbench.py cfg3 --graph --only warp > gpurun_out/c15_kb_staged$st.log 2>&1
done
tail -n 5 gpurun_out/c15_pytest.log; grep -h -i warp gpurun_out/c15_kb_staged*.log
