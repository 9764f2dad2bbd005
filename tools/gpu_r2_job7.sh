#!/bin/bash
O=gpurun_out
( time python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "ntt" ) > $O/r2_t7.log 2>&1; tail -5 $O/r2_t7.log
python bench_extra.py ntt64 > $O/r2_ntt64_v1.log 2>&1; cat $O/r2_ntt64_v1.log | cut -c1-700
