#!/bin/bash
# r2 job 11: the allocation-free entry point alone in a fresh process, then the whole -m gpu suite with the final build,
# then the 1-GPU bench and smoke()
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > gpurun_out/j11_smi.txt
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q > gpurun_out/j11_noalloc_alone.log 2>&1; echo "alone rc=$?" >> gpurun_out/j11_noalloc_alone.log
timeout 3000 python -m pytest tests -m gpu -q --durations=15 > gpurun_out/j11_pytest_full.log 2>&1; echo "full rc=$?" >> gpurun_out/j11_pytest_full.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/j11_smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/j11_smoke.log
timeout 900 python bench.py --steps 3 --warmup 3 > gpurun_out/j11_bench.json 2> gpurun_out/j11_bench.err; echo "bench rc=$?" >> gpurun_out/j11_bench.err
tail -3 gpurun_out/j11_noalloc_alone.log gpurun_out/j11_pytest_full.log gpurun_out/j11_smoke.log gpurun_out/j11_bench.err
