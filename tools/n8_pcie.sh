#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29521 tools/pcie_aggregate.py 2> gpurun_out/pcie_n8.err | grep n_gpus > gpurun_out/pcie_n8.json; cat gpurun_out/pcie_n8.json
timeout 120 python tools/pcie_aggregate.py 2> gpurun_out/pcie_n1.err | grep n_gpus > gpurun_out/pcie_n1.json; cat gpurun_out/pcie_n1.json
