#!/bin/bash
# round-end evidence on one B200: full GPU parity suite, smoke, the default bench line, the ncu launch list of the same
# command, and one ncu --set full capture of the K1 kernels (each ncu pass only after the plain run exited 0)
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_full.log 2>&1; tail -3 gpurun_out/pytest_gpu_full.log
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
python bench.py > gpurun_out/bench_final.json 2> gpurun_out/bench_final.err || exit 1
python -c "
import json; d=json.load(open('gpurun_out/bench_final.json')); print(d['value'], d['ms_per_step'], d['e2e']['value'], d['stages_ms'], d['roofline']['frac'], d['cpu_baseline']['value'], d['clocks'])"
python bench.py --steps 2 --warmup 3 --no-cpu-baseline > /dev/null 2> gpurun_out/pre.err || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_final.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_launches.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:"pipe|k_decode_raw" -s 24 -c 6 -o gpurun_out/r01_ncu_full_v5_k1 -f python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_k1.log 2>&1
ls -la gpurun_out/*.ncu-rep gpurun_out/launches_final.csv
python tools/microbench_dtw.py 2>&1 | tail -3 | cut -c1-300
