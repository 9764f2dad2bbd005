#!/bin/bash
# r2 job 38: BGV branch of ApproxModDown on the device vs the oracle; RNS + static-scratch tests 3 times
O=gpurun_out; mkdir -p $O
for i in 1 2 3; do timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_zz_no_alloc.py -m gpu -q -k "rns or static" 2>&1 | tail -n 1; done > $O/j38_rns.log; cat $O/j38_rns.log
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -q -k "rns" 2>&1 | tail -n 15 | cut -c1-200 > $O/j38_detail.log; tail -n 4 $O/j38_detail.log
