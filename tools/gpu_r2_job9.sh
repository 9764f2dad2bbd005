#!/bin/bash
O=gpurun_out
( time python -m pytest tests/test_gpu_parity.py tests/test_host_cpp.py -m gpu -q -x -k "btkeygen or device_keygen" --durations=8 ) > $O/r2_t9.log 2>&1; tail -14 $O/r2_t9.log
