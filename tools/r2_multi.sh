#!/bin/bash
# usage: r2_multi.sh N   (inside gpurun --gpus N)
N=$1
mkdir -p gpurun_out/m$N
nvidia-smi topo -m > gpurun_out/m$N/topo.txt 2>&1
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 tools/pcie_probe_ranks.py > gpurun_out/m$N/pcie.log 2>&1
grep pcie_probe gpurun_out/m$N/pcie.log
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29518 bench.py --gpus $N --steps 100 --warmup 5 > gpurun_out/m$N/bench.json 2> gpurun_out/m$N/bench.err
cat gpurun_out/m$N/bench.json
tail -n 5 gpurun_out/m$N/bench.err
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29519 bench.py --impl reference --gpus $N --steps 3 --warmup 1 > gpurun_out/m$N/bench_ref.json 2> gpurun_out/m$N/bench_ref.err
cat gpurun_out/m$N/bench_ref.json
