#!/bin/bash
# A/B of prebuilt library variants (ab/*.so) on the default bench (C2 map loop + dense hot path), two rounds interleaved.
mkdir -p gpurun_out
cp rawalign_b200/librawalign_b200.so /tmp/lib_default.so
for round in 1 2; do
for so in ab/*.so; do
  cp "$so" rawalign_b200/librawalign_b200.so
  python bench.py --steps ${AB_STEPS:-12} --warmup 3 --no-cpu-baseline 2> gpurun_out/ab.err | python -c "
import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); h=d['hot_path']; print('$so', 'r$round', 'map %.0f reads/s (%.2f ms; alone: dtw %.2f of %.2f ms)' % (d['value'], d['ms_per_step'], d['map_stages_ms']['dtw_ms'], d['map_stages_ms']['total_ms']), 'e2e %.0f' % d['e2e']['value'], '| dense %.0f reads/s, chain dtw %.3f ms = %.0f GCUPS' % (h['value'], h['stages_ms']['chain_dtw_kernels'], h['dtw_gcups']))" | tee -a gpurun_out/ab_map.log
done
done
cp /tmp/lib_default.so rawalign_b200/librawalign_b200.so
