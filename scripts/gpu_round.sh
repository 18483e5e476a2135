#!/bin/bash
# One gpurun call: GPU parity tests (the tcgen05 tests first, in their own process, so a fault there cannot poison the
# rest), smoke, bench, then the ncu launch list and one full capture of the search kernel (each ncu run directly after
# the same command exited 0 without ncu).  Logs land in gpurun_out/.   usage: gpu_round.sh [noncu]
set -u
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt 2>&1
timeout 900 python -m pytest tests/test_gpu_search_tc.py tests/test_gpu_search_prepared.py -q -m gpu -p no:cacheprovider > gpurun_out/pytest_tc.log 2>&1
TC=$?; echo "pytest tc exit $TC" >> gpurun_out/pytest_tc.log; tail -n 25 gpurun_out/pytest_tc.log
if [ $TC -ne 0 ]; then export LIPVQ_SEARCH_ALGO=simt; echo "tcgen05 tests failed: rest of the round runs with LIPVQ_SEARCH_ALGO=simt"; fi
timeout 900 python -m pytest tests -q -m gpu -p no:cacheprovider --deselect tests/test_gpu_search_tc.py --deselect tests/test_gpu_search_prepared.py > gpurun_out/pytest_gpu.log 2>&1
echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -n 15 gpurun_out/pytest_gpu.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke exit $?"; tail -n 3 gpurun_out/smoke.log
timeout 600 python bench.py --steps 30 --warmup 5 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit $?"
cat gpurun_out/bench.json; tail -n 5 gpurun_out/bench.err
if [ "${1:-}" != "noncu" ]; then
  timeout 300 python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-t2 > gpurun_out/plain.log 2>&1 &&
  timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches.csv \
      python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-t2 > gpurun_out/ncu_launches.log 2>&1
  echo "ncu launches exit $?"
  timeout 300 python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-t2 > gpurun_out/plain2.log 2>&1 &&
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:"${NCU_KERNEL:-search_p1}" -s 6 -c 3 \
      -f -o gpurun_out/prof_search python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-t2 > gpurun_out/ncu_full.log 2>&1
  echo "ncu full exit $?"
fi
ls -la gpurun_out | head -30
