#!/bin/bash
# lookup timeline (bring-up build made on the box only; the tree's library stays the shipped one)
mkdir -p gpurun_out; rm -f gpurun_out/c26*
touch cardiac-segmentation-optical-flow_b200/csrc/lookup.cu cardiac-segmentation-optical-flow_b200/csrc/cabi.cu cardiac-segmentation-optical-flow_b200/csrc/corr_tcgen05.cu
CF_NVCC_EXTRA=-DCF_BRINGUP timeout 900 python cardiac-segmentation-optical-flow_b200/_build.py > gpurun_out/c26_build.log 2>&1 || { tail gpurun_out/c26_build.log; exit 1; }
timeout 300 python tools/lookup_ts.py cfg2 > gpurun_out/c26_ts_cfg2.log 2>&1; cat gpurun_out/c26_ts_cfg2.log
timeout 300 python tools/lookup_ts.py cfg4 > gpurun_out/c26_ts_cfg4.log 2>&1; cat gpurun_out/c26_ts_cfg4.log
