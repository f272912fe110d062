#!/bin/bash
# parity of the chain paths, then A/B of the sparse part kernels (warp tiles vs CTA tiles) on one box, then a launch list
mkdir -p gpurun_out
python tools/dbg_sp2.py 2>&1 | tail -14
python -m pytest tests -m gpu -x -q -k "chain or sparse" > gpurun_out/pytest_gpu.log 2>&1; tail -5 gpurun_out/pytest_gpu.log
for r in 0 8; do
RA_BENCH_RESERVED=$r python bench.py --steps 10 --warmup 3 --no-cpu-baseline 2> gpurun_out/bench.err | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('reserved=$r', 'value %.0f' % d['value'], 'ms %.3f' % d['ms_per_step'], 'K1 %.3f' % d['stages_ms']['events_all_kernels'], 'dtw %.3f' % d['stages_ms']['chain_dtw_kernels'], 'gcups %.0f' % d['dtw_gcups'], 'e2e %.0f' % d['e2e']['value'])" | tee -a gpurun_out/sp2.log
done
ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum,smsp__thread_inst_executed_per_inst_executed.ratio,smsp__issue_active.avg.pct_of_peak_sustained_active --clock-control none -k regex:"k_sp_" -s 12 -c 8 --csv --log-file gpurun_out/launches_sp2.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_bench.log 2>&1
python - <<'PY'
import csv
rows=[r for r in csv.reader(open('gpurun_out/launches_sp2.csv')) if len(r)>5]
h=rows[0]; ki=h.index('Kernel Name'); mi=h.index('Metric Name'); vi=h.index('Metric Value'); ii=h.index('ID')
cur={}
for r in rows[1:]:
    cur.setdefault((r[ii],r[ki][:30]),{})[r[mi].split('.')[0][-22:]]=r[vi]
for k,v in cur.items(): print(k, v)
PY
