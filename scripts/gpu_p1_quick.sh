#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_search_prepared.py -x -q -m gpu -p no:cacheprovider 2>&1 | tail -3
for mode in t1 uniform; do
  echo "== prepared $mode"; timeout 120 python scripts/search_once.py 65536 512 64 $mode 5 --prepared 2>&1 | tail -3
done
echo "== ties 437"; timeout 120 python scripts/search_once.py 65536 512 64 t1 5 --prepared --ties=437 2>&1 | tail -3
echo "== ties 3000"; timeout 120 python scripts/search_once.py 65536 512 64 t1 5 --prepared --ties=3000 2>&1 | tail -3
timeout 600 python bench.py --steps 30 --warmup 5 --no-cpu-baseline --no-sweep > gpurun_out/bench_q.json 2> gpurun_out/bench_q.err; echo "bench exit $?"; tail -3 gpurun_out/bench_q.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/bench_q.json'))
print(d['ms_per_step'], d['value'], d['e2e']['value'], d['tokenize']['value'], d['tokenize']['ms_per_step'], d['tokenize_e2e']['value'])
r=d['roofline']
for k in ('us_per_launch','us_per_launch_isolated','us_per_launch_writeflush','frac','us_prepare','us_unprepared','rows_uncertified_by_fp16_pass'): print(k, r.get(k))
PY
