#!/bin/bash
# r2 job 45: ncu --set full of the RNS arithmetic kernels inside one hybrid key switch batch
O=gpurun_out; mkdir -p $O
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"rns_inner|rns_switch" -c 3 -o $O/j45_rns python tools/prof_hks.py 1 > $O/j45_ncu.log 2>&1; echo "ncu rc=$?" >> $O/j45_ncu.log; tail -n 1 $O/j45_ncu.log
