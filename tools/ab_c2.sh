#!/bin/bash
# A/B of prebuilt library variants (ab/*.so) on the C2 map loop alone, two rounds interleaved.
mkdir -p gpurun_out
cp rawalign_b200/librawalign_b200.so /tmp/lib_default.so
for round in 1 2; do
for so in ab/*.so; do
  cp "$so" rawalign_b200/librawalign_b200.so
  python bench.py --config ${AB_CONFIG:-C2} --steps ${AB_STEPS:-16} --warmup 3 --no-hot-path --no-cpu-baseline 2> gpurun_out/ab.err | python -c "
import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); s=d['map_stages_ms']; print('$so', 'r$round', 'map %.0f reads/s (%.2f ms)' % (d['value'], d['ms_per_step']), 'e2e %.0f' % d['e2e']['value'], 'alone: chain %.2f sort %.2f dtw %.2f events %.2f total %.2f' % (s['chain_ms'], s['sort_ms'], s['dtw_ms'], s['events_ms'], s['total_ms']))" | tee -a gpurun_out/ab_c2.log
done
done
cp /tmp/lib_default.so rawalign_b200/librawalign_b200.so
