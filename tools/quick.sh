#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q -k "events" 2>&1 | tail -1
for i in 1 2 3; do
python bench.py --steps 12 --warmup 3 --no-cpu-baseline 2> gpurun_out/bench.err | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('run $i', 'value %.0f' % d['value'], 'ms %.3f' % d['ms_per_step'], 'K1 %.3f' % d['stages_ms']['events_all_kernels'], 'dtw %.3f' % d['stages_ms']['chain_dtw_kernels'], 'e2e %.0f' % d['e2e']['value'])"
done
