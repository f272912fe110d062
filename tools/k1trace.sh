#!/bin/bash
mkdir -p gpurun_out
RA_BENCH_PIPELINE=1 RA_K1_TRACE=1 python bench.py --steps 3 --warmup 3 --no-cpu-baseline 2> gpurun_out/k1trace.err > /dev/null
grep "k1 trace" gpurun_out/k1trace.err | tail -16
