#!/bin/bash
# r2 job 24: two-word reduction in the quad kernels, default-stream copy races closed: the tests that failed + the N = 2048 /
# RNS / keygen groups, each RNS case 5 times in a row; then throughput
O=gpurun_out; mkdir -p $O
timeout 1800 python -m pytest tests/test_gpu_parity.py -m gpu -q -k "wide or keygen or rns" --durations=5 > $O/j24_pytest.log 2>&1; echo "pytest rc=$?" >> $O/j24_pytest.log
tail -n 12 $O/j24_pytest.log | cut -c1-250
for i in 1 2 3 4 5; do timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_zz_no_alloc.py -m gpu -q -k "rns or static" 2>&1 | tail -n 1; done > $O/j24_rns_repeat.log; cat $O/j24_rns_repeat.log
timeout 900 python tools/bench_wide.py 1776 > $O/j24_wide_bench.jsonl 2> $O/j24_wide_bench.err; echo "rc=$?" >> $O/j24_wide_bench.err
head -n 2 $O/j24_wide_bench.jsonl | cut -c1-220
