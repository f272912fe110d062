#!/bin/bash
# 8-GPU run of bench.py with topology notes (host sockets, GPU<->CPU affinity) written next to the bench line.
mkdir -p gpurun_out
{ nvidia-smi topo -m; lscpu | grep -iE "numa|socket|model name|^CPU\(s\)"; nproc; free -g | head -2; } > gpurun_out/n8_topo.txt 2>&1
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 8 --steps 6 --warmup 3 > gpurun_out/bench_n8_bind.json 2> gpurun_out/bench_n8_bind.err
echo rc=$?; cat gpurun_out/bench_n8_bind.json | head -c 3000
