#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q -k "events" > gpurun_out/pytest_gpu.log 2>&1; tail -2 gpurun_out/pytest_gpu.log
for cfg in "8 1" "12 1" "16 1" "8 1" "12 1" "16 1"; do set -- $cfg
RA_K1_GROUPS=$1 RA_K1_PRIO=$2 python bench.py --steps 12 --warmup 3 --no-cpu-baseline 2> gpurun_out/bench.err | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('groups=$1 prio=$2', 'value %.0f' % d['value'], 'ms %.3f' % d['ms_per_step'], 'K1 %.3f' % d['stages_ms']['events_all_kernels'], 'dtw %.3f' % d['stages_ms']['chain_dtw_kernels'], 'e2e %.0f (%.2f ms)' % (d['e2e']['value'], d['e2e']['ms_per_step']))" | tee -a gpurun_out/ab3.log
done
