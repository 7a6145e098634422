#!/bin/bash
# refresh of the logs that the last kernel changes touch (bounded)
O=gpurun_out/r2; mkdir -p $O
timeout 600 python -m pytest tests -q -x -m gpu > $O/r2_pytest_gpu.log 2>&1; echo "rc=$?" >> $O/r2_pytest_gpu.log; tail -n 3 $O/r2_pytest_gpu.log
timeout 200 python __graft_entry__.py --smoke > $O/r2_smoke.log 2>&1; tail -n 1 $O/r2_smoke.log
for c in cfg2 cfg4; do timeout 300 python tools/kbench.py $c --graph > $O/r2_kbench_$c.log 2>&1; grep -i "backward" $O/r2_kbench_$c.log | cut -c1-110; done
