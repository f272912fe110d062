#!/bin/bash
# A/B of prebuilt library variants (ab/*.so, built here with `make EXTRA=-D...`): each is copied over the in-tree library and
# benched on the same box, two rounds interleaved, so box-to-box variation does not blur the comparison.
mkdir -p gpurun_out
cp rawalign_b200/librawalign_b200.so /tmp/lib_default.so
for round in 1 2; do
for so in ab/*.so; do
  cp "$so" rawalign_b200/librawalign_b200.so
  python bench.py --steps ${AB_STEPS:-10} --warmup 3 --no-cpu-baseline 2> gpurun_out/ab.err | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('$so', 'r$round', 'value %.0f' % d['value'], 'ms %.3f' % d['ms_per_step'], 'K1 %.3f' % d['stages_ms']['events_all_kernels'], 'dtw %.3f' % d['stages_ms']['chain_dtw_kernels'], 'e2e %.0f (%.2f ms)' % (d['e2e']['value'], d['e2e']['ms_per_step']))" | tee -a gpurun_out/ab.log
done
done
cp /tmp/lib_default.so rawalign_b200/librawalign_b200.so
