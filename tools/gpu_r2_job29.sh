#!/bin/bash
# r2 job 29: hunt the intermittent crash of `host_parity TOY_LMKCDEY 8 1 --gpu-keygen` (1 of 4 full-suite runs): 60 runs with a
# backtrace handler; then the N = 2048 tests and throughput with the vectorised cross-stage exchange
O=gpurun_out; mkdir -p $O
: > $O/j29_crash.log
fails=0
for i in $(seq 1 60); do
  timeout 120 host/_build/host_parity TOY_LMKCDEY 8 1 --gpu-keygen > $O/j29_run.out 2> $O/j29_run.err; rc=$?
  if [ $rc -ne 0 ]; then fails=$((fails+1)); echo "=== run $i rc=$rc" >> $O/j29_crash.log; tail -n 5 $O/j29_run.out >> $O/j29_crash.log; cat $O/j29_run.err >> $O/j29_crash.log; fi
done
echo "TOY_LMKCDEY --gpu-keygen: $fails failures of 60" | tee -a $O/j29_crash.log
fails=0
for i in $(seq 1 30); do
  timeout 120 host/_build/host_parity TOY 8 1 --gpu-keygen > $O/j29_run.out 2> $O/j29_run.err; rc=$?
  if [ $rc -ne 0 ]; then fails=$((fails+1)); echo "=== TOY run $i rc=$rc" >> $O/j29_crash.log; tail -n 5 $O/j29_run.out >> $O/j29_crash.log; cat $O/j29_run.err >> $O/j29_crash.log; fi
done
echo "TOY --gpu-keygen: $fails failures of 30" | tee -a $O/j29_crash.log
head -n 60 $O/j29_crash.log | cut -c1-200
timeout 1800 python -m pytest tests/test_gpu_parity.py -m gpu -q -k "wide or keygen" > $O/j29_pytest.log 2>&1; echo "pytest rc=$?" >> $O/j29_pytest.log
tail -n 3 $O/j29_pytest.log | cut -c1-250
timeout 900 python tools/bench_wide.py 1776 > $O/j29_wide_bench.jsonl 2> $O/j29_wide_bench.err; echo "rc=$?" >> $O/j29_wide_bench.err
grep "STD256" $O/j29_wide_bench.jsonl | cut -c1-220
