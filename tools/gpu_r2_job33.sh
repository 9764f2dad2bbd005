#!/bin/bash
# r2 job 33: A/B — key prefetch (prefetch.global.L2 at round start) in br_sched_kernel
O=gpurun_out; mkdir -p $O
for v in main pf main pf; do
  if [ $v = main ]; then so=fhe_b200/libluxb200.so; else so=fhe_b200/_dev/libluxb200_$v.so; fi
  LUXB200_SO=$PWD/$so timeout 600 python tools/ab_sched.py 2>&1 | grep "^{" | tee -a $O/j33_ab.jsonl
done
