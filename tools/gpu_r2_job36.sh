#!/bin/bash
# r2 job 36: throughput of the three-digit N = 1024 sets (br_ginx_kernel<10,..,6>, 4 ciphertexts per SM, keys through L2)
O=gpurun_out; mkdir -p $O
timeout 600 python tools/bench_wide.py 4736 --n1024 > $O/j36_n1024_rows6.jsonl 2> $O/j36.err; echo "rc=$?" >> $O/j36.err
cut -c1-230 $O/j36_n1024_rows6.jsonl; tail -n 2 $O/j36.err
