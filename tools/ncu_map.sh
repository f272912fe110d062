#!/bin/bash
# ncu --set full of the map loop's own kernels (one launch each of the first, full-size round) -> gpurun_out/*.ncu-rep
K=${1:-k_chain}
RA_BENCH_MAP_DEPTH=1 ncu --set full --clock-control none --import-source on -k regex:$K -c 1 -o gpurun_out/r02_ncu_$K -f \
   python bench.py --steps 1 --warmup 3 --no-hot-path --no-cpu-baseline > gpurun_out/ncu_full_$K.log 2>&1
tail -2 gpurun_out/ncu_full_$K.log | cut -c1-200
