#!/bin/bash
# r2 job 44: kernel-time breakdown of one hybrid key switch batch (ncu launch list)
O=gpurun_out; mkdir -p $O
timeout 600 python tools/prof_hks.py 2 > $O/j44_plain.log 2>&1; echo "plain rc=$?" >> $O/j44_plain.log; tail -n 1 $O/j44_plain.log
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file $O/j44_hks_launches.csv python tools/prof_hks.py 2 > $O/j44_ncu.log 2>&1; echo "ncu rc=$?" >> $O/j44_ncu.log; tail -n 1 $O/j44_ncu.log
wc -l $O/j44_hks_launches.csv
