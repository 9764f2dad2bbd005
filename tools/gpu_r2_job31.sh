#!/bin/bash
# r2 job 31: four-digit N = 2048 sets on the four-warp kernel: parity and throughput
O=gpurun_out; mkdir -p $O
timeout 1800 python -m pytest tests/test_gpu_parity.py -m gpu -q -k "wide or keygen" --durations=4 > $O/j31_pytest.log 2>&1; echo "pytest rc=$?" >> $O/j31_pytest.log
tail -n 9 $O/j31_pytest.log | cut -c1-250
timeout 900 python tools/bench_wide.py 1776 > $O/j31_wide_bench.jsonl 2> $O/j31_wide_bench.err; echo "rc=$?" >> $O/j31_wide_bench.err
grep "STD256" $O/j31_wide_bench.jsonl | cut -c1-220
