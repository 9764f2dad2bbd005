#!/bin/bash
# r2 job 22: quad kernel on hand-made N = 2048 sets x {GINX, AP, LMKCDEY}, then its ncu profile
O=gpurun_out; mkdir -p $O
timeout 1500 python -m pytest tests/test_gpu_parity.py -m gpu -q -k "wide_custom" --durations=5 > $O/j22_pytest.log 2>&1; echo "pytest rc=$?" >> $O/j22_pytest.log
tail -n 25 $O/j22_pytest.log | cut -c1-300
timeout 900 ncu --set full --clock-control none --import-source on -k regex:br_quad_kernel -c 1 -o $O/j22_quad python tools/prof_wide.py STD256 444 > $O/j22_ncu.log 2>&1; echo "ncu rc=$?" >> $O/j22_ncu.log
tail -n 3 $O/j22_ncu.log
