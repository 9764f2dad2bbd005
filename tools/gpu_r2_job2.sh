#!/bin/bash
O=gpurun_out
( time python -m pytest tests -m gpu -q --durations=12 ) > $O/r2_t2.log 2>&1; tail -40 $O/r2_t2.log
