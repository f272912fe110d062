#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_full.log 2>&1; tail -2 gpurun_out/pytest_gpu_full.log
cp rawalign_b200/librawalign_b200.so /tmp/keep.so
AB_STEPS=12 bash tools/ab.sh
cp /tmp/keep.so rawalign_b200/librawalign_b200.so
# K1 latency of a small batch (512 reads): sequential state machine vs four segments
for seg in 1 4; do
RA_FSM_SEG=$seg RA_BENCH_PIPELINE=1 python bench.py --reads-per-gpu 512 --steps 10 --warmup 3 --no-cpu-baseline 2> gpurun_out/bench.err | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('512 reads, one batch in flight, seg=$seg', 'value %.0f' % d['value'], 'ms %.3f' % d['ms_per_step'], 'K1 %.3f' % d['stages_ms']['events_all_kernels'], 'dtw %.3f' % d['stages_ms']['chain_dtw_kernels'])" | tee -a gpurun_out/ab.log
done
