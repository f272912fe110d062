#!/bin/bash
mkdir -p gpurun_out
bash tools/ab.sh
for d in 1 2 3 4; do
RA_BENCH_PIPELINE=$d python bench.py --steps 12 --warmup 3 --no-cpu-baseline 2> gpurun_out/bench.err | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('depth=$d', 'value %.0f' % d['value'], 'ms %.3f' % d['ms_per_step'], 'e2e %.0f (%.2f ms)' % (d['e2e']['value'], d['e2e']['ms_per_step']))" | tee -a gpurun_out/ab.log
done
