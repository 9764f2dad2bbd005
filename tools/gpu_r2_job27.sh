#!/bin/bash
# r2 job 27: LMKCDEY at N = 2048 on the four-warp kernel: oracle parity (named + hand-made sets), live reference, throughput
O=gpurun_out; mkdir -p $O
timeout 1800 python -m pytest tests/test_gpu_parity.py -m gpu -q -k "wide or keygen" --durations=5 > $O/j27_pytest.log 2>&1; echo "pytest rc=$?" >> $O/j27_pytest.log
tail -n 14 $O/j27_pytest.log | cut -c1-250
timeout 1500 python -m pytest tests/test_host_cpp.py -m gpu -q -k "STD256_LMKCDEY" > $O/j27_host.log 2>&1; echo "host rc=$?" >> $O/j27_host.log; tail -n 3 $O/j27_host.log
timeout 900 python tools/bench_wide.py 1776 > $O/j27_wide_bench.jsonl 2> $O/j27_wide_bench.err; echo "rc=$?" >> $O/j27_wide_bench.err
cut -c1-220 $O/j27_wide_bench.jsonl
