#!/bin/bash
mkdir -p gpurun_out
bash tools/k1trace.sh | tail -8
AB_STEPS=12 bash tools/ab.sh
cp ab/lib_b_fsmseg.so rawalign_b200/librawalign_b200.so
for seg in 1 2 4; do
RA_FSM_SEG=$seg python bench.py --steps 12 --warmup 3 --no-cpu-baseline 2> gpurun_out/bench.err | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('seg=$seg', 'value %.0f' % d['value'], 'ms %.3f' % d['ms_per_step'], 'K1 %.3f' % d['stages_ms']['events_all_kernels'], 'dtw %.3f' % d['stages_ms']['chain_dtw_kernels'])" | tee -a gpurun_out/ab.log
done
