#!/bin/bash
O=gpurun_out
( time python -m pytest tests/test_circuit.py tests/test_host_cpp.py -m gpu -q -x --durations=5 ) > $O/r2_t5.log 2>&1; tail -30 $O/r2_t5.log
