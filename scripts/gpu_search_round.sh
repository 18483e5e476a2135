#!/bin/bash
# search-only GPU round: parity tests of the nearest-code search, the C3 sweep, and the per-kernel ncu launch list
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_search_tc.py -q -m gpu -p no:cacheprovider 2>&1 | tail -4 | tee gpurun_out/pytest_tc.log
timeout 300 python -m pytest tests/test_gpu_kernels.py -q -m gpu -p no:cacheprovider 2>&1 | tail -5 | tee gpurun_out/pytest_k.log
timeout 600 python scripts/bench_search.py ${1:---quick} 2>&1 | tee gpurun_out/sweep.log
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/sweep_launches.csv python scripts/bench_search.py --quick --profile > gpurun_out/sweep_ncu.log 2>&1
