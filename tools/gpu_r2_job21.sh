#!/bin/bash
# r2 job 21: br_quad.cuh (N = 2048 as four warps per ciphertext): parity of the N = 2048 sets, keygen, throughput
O=gpurun_out; mkdir -p $O
timeout 1500 python -m pytest tests/test_gpu_parity.py -m gpu -q -k "wide or keygen" --durations=5 > $O/j21_pytest.log 2>&1; echo "pytest rc=$?" >> $O/j21_pytest.log
tail -n 25 $O/j21_pytest.log | cut -c1-300
timeout 900 python tools/bench_wide.py 1184 > $O/j21_wide_bench.jsonl 2> $O/j21_wide_bench.err; echo "rc=$?" >> $O/j21_wide_bench.err
cut -c1-220 $O/j21_wide_bench.jsonl; tail -n 2 $O/j21_wide_bench.err
