#!/bin/bash
O=gpurun_out
( time python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "btkeygen" --durations=8 ) > $O/r2_t8.log 2>&1; tail -30 $O/r2_t8.log
