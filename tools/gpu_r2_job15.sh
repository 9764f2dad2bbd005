#!/bin/bash
# r2 job 15: the N = 2048 two-digit sets on the register-resident kernels (br11.cu): parity tests that touch them, throughput
O=gpurun_out; mkdir -p $O
timeout 1500 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "wide or keygen or keyswitch_tile or std128_full" --durations=8 > $O/j15_pytest.log 2>&1; echo "pytest rc=$?" >> $O/j15_pytest.log
tail -n 14 $O/j15_pytest.log
timeout 900 python tools/bench_wide.py 1184 > $O/j15_wide_bench.jsonl 2> $O/j15_wide_bench.err; echo "rc=$?" >> $O/j15_wide_bench.err
cat $O/j15_wide_bench.jsonl | cut -c1-250; tail -n 3 $O/j15_wide_bench.err
