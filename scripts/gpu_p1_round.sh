#!/bin/bash
# GPU round for the prepared / in-kernel-tail search: parity tests first, then timings of the two entry points
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_search_prepared.py -x -q -m gpu -p no:cacheprovider 2>&1 | tail -15 | tee gpurun_out/pytest_prep.log
timeout 600 python -m pytest tests/test_gpu_search_tc.py -x -q -m gpu -p no:cacheprovider 2>&1 | tail -8 | tee gpurun_out/pytest_tc.log
for mode in t1 uniform; do
  echo "== unprepared $mode"; timeout 120 python scripts/search_once.py 65536 512 64 $mode 6 2>&1 | tail -5
  echo "== prepared $mode"; timeout 120 python scripts/search_once.py 65536 512 64 $mode 6 --prepared 2>&1 | tail -5
done
echo "== prepared t1 N=1M"; timeout 120 python scripts/search_once.py 1048576 512 64 t1 6 --prepared 2>&1 | tail -4
echo "== ties 437"; timeout 120 python scripts/search_once.py 65536 512 64 t1 6 --prepared --ties=437 2>&1 | tail -4
echo "== ties 3000"; timeout 120 python scripts/search_once.py 65536 512 64 t1 6 --prepared --ties=3000 2>&1 | tail -4
echo "== trace prepared cold"; timeout 120 python scripts/trace_search.py 65536 512 64 t1 --prepared --cold 2>&1 | tee gpurun_out/trace_prepared_cold.txt | tail -70
if [ "${1:-}" = "full" ]; then
timeout 900 python -m pytest tests -x -q -m gpu -p no:cacheprovider --deselect tests/test_gpu_search_tc.py --deselect tests/test_gpu_search_prepared.py 2>&1 | tail -8 | tee gpurun_out/pytest_rest.log
timeout 600 python bench.py --steps 30 --warmup 5 --no-cpu-baseline > gpurun_out/bench_p1.json 2> gpurun_out/bench_p1.err; echo "bench exit $?"; cat gpurun_out/bench_p1.json | cut -c1-3000; tail -3 gpurun_out/bench_p1.err
fi
