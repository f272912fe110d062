#!/bin/bash
# 8-GPU evidence: aggregate host<->device bandwidth with all ranks copying at once (the ceiling of e2e at N=8), then the bench
mkdir -p gpurun_out
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29521 tools/pcie_aggregate.py > gpurun_out/pcie_n8.json 2> gpurun_out/pcie_n8.err; echo pcie rc=$?; cat gpurun_out/pcie_n8.json
timeout 120 python tools/pcie_aggregate.py > gpurun_out/pcie_n1.json 2> gpurun_out/pcie_n1.err; cat gpurun_out/pcie_n1.json
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 8 --steps 10 --warmup 3 > gpurun_out/bench_n8_final.json 2> gpurun_out/bench_n8_final.err
echo bench rc=$?; python -c "
import json; d=json.load(open('gpurun_out/bench_n8_final.json')); print(d['value'], d['ms_per_step'], d['e2e'])"
