#!/bin/bash
mkdir -p gpurun_out/c16
timeout 200 python tools/aten_cuda_probe.py > gpurun_out/c16/probe.log 2>&1
cat gpurun_out/c16/probe.log
