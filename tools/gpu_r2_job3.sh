#!/bin/bash
O=gpurun_out
( time python -m pytest tests -m gpu -q -x -k "largeprec or dcrtpoly or e2e_cpp" --durations=5 ) > $O/r2_t3.log 2>&1; tail -30 $O/r2_t3.log
python tools/bench_wide.py 1184 > $O/r2_wide_bench.log 2>&1; cat $O/r2_wide_bench.log | cut -c1-300
