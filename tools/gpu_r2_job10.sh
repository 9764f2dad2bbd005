#!/bin/bash
O=gpurun_out
( time python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "rns" --durations=6 ) > $O/r2_t10.log 2>&1; tail -14 $O/r2_t10.log
