#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q -k "dtw" > gpurun_out/pytest_gpu.log 2>&1; tail -3 gpurun_out/pytest_gpu.log
python tools/microbench_dtw.py 2>&1 | tail -9
