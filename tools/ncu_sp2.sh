#!/bin/bash
mkdir -p gpurun_out
python bench.py --steps 2 --warmup 3 --no-cpu-baseline > /dev/null 2> gpurun_out/pre.err || exit 1
ncu --set full --clock-control none --import-source on -k regex:"k_sp_parts2|k_sp_rest" -s 4 -c 2 -o gpurun_out/r01_ncu_full_v5_sp2 -f python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_sp2.log 2>&1
ls -la gpurun_out/*.ncu-rep
