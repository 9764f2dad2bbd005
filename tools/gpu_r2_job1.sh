#!/bin/bash
# round-2 GPU job 1: full GPU test suite, bench line, C++ e2e, launch list + DRAM traffic of the bench command, racecheck
O=gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm --format=csv,noheader > $O/r2_gpu.txt
nproc >> $O/r2_gpu.txt
( time python -m pytest tests -m gpu -x -q --durations=15 ) > $O/r2_t1.log 2>&1; tail -25 $O/r2_t1.log
python bench.py --steps 3 --warmup 3 > $O/r2_bench1.log 2>&1; tail -1 $O/r2_bench1.log | cut -c1-600
host/_build/host_parity STD128 6 0 --e2e 65536 > $O/r2_e2e_cpp.log 2>&1; grep e2e_cpp $O/r2_e2e_cpp.log
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 400 --csv --log-file $O/r2_launches_v1.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-extras --no-strong > $O/r2_ncu_bench.log 2>&1; wc -l $O/r2_launches_v1.csv
for w in GINX LMKCDEY; do timeout 600 compute-sanitizer --tool racecheck python tools/sanitize_toy.py $w > $O/r2_racecheck_$w.log 2>&1; tail -4 $O/r2_racecheck_$w.log; done
