python -m pytest tests -m gpu -x -q -k "events" > gpurun_out/pytest_gpu.log 2>&1; tail -3 gpurun_out/pytest_gpu.log
python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench.json 2> gpurun_out/bench.err && python -c "
import json; d=json.load(open('gpurun_out/bench.json')); print(d['value'], d['stages_ms'], d['e2e']['ms_per_step'])"
ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum,smsp__thread_inst_executed_per_inst_executed.ratio,smsp__issue_active.avg.pct_of_peak_sustained_active --clock-control none -k regex:"pipe|scan" -s 20 -c 5 --csv --log-file gpurun_out/launches_v4.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_bench.log 2>&1
python - <<'PY'
import csv
rows=[r for r in csv.reader(open('gpurun_out/launches_v4.csv')) if len(r)>5]
h=rows[0]; ki=h.index('Kernel Name'); mi=h.index('Metric Name'); vi=h.index('Metric Value'); ii=h.index('ID')
cur={}
for r in rows[1:]:
    cur.setdefault((r[ii],r[ki][:30]),{})[r[mi].split('.')[0][-22:]]=r[vi]
for k,v in cur.items(): print(k, v)
PY
