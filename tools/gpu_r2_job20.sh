#!/bin/bash
# r2 job 20: brw_kernel<u32> capped at 64 registers (4 CTAs per SM with 2 digits): parity of the wide sets, throughput
O=gpurun_out; mkdir -p $O
timeout 1500 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "wide or keygen" > $O/j20_pytest.log 2>&1; echo "pytest rc=$?" >> $O/j20_pytest.log
tail -n 3 $O/j20_pytest.log
timeout 900 python tools/bench_wide.py 1184 > $O/j20_wide_bench.jsonl 2> $O/j20_wide_bench.err; echo "rc=$?" >> $O/j20_wide_bench.err
cut -c1-220 $O/j20_wide_bench.jsonl; tail -n 2 $O/j20_wide_bench.err
