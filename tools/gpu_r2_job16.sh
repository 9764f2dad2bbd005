#!/bin/bash
# r2 job 16: ncu of the N = 2048 register-resident kernels
O=gpurun_out; mkdir -p $O
timeout 900 ncu --set full --clock-control none --import-source on -k regex:br_ginx_kernel -c 1 -o $O/j16_ginx11 python tools/prof_wide.py STD256 296 > $O/j16_ncu_ginx.log 2>&1; echo "ncu rc=$?" >> $O/j16_ncu_ginx.log
timeout 900 ncu --set full --clock-control none --import-source on -k regex:br_sched_kernel -c 1 -o $O/j16_sched11 python tools/prof_wide.py STD256_LMKCDEY 296 > $O/j16_ncu_sched.log 2>&1; echo "ncu rc=$?" >> $O/j16_ncu_sched.log
tail -n 2 $O/j16_ncu_ginx.log $O/j16_ncu_sched.log
