#!/bin/bash
# r2 job 13: key-ordered key switch (ks_tile.cuh): parity, bench with it on / off, ncu of the two new kernels
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "keyswitch_tile or std128_full_batch or golden" --durations=5 > $O/j13_pytest.log 2>&1; echo "pytest rc=$?" >> $O/j13_pytest.log
tail -4 $O/j13_pytest.log
grep -q "pytest rc=0" $O/j13_pytest.log || exit 1
timeout 600 python bench.py --steps 3 --warmup 3 --no-extras --no-strong > $O/j13_bench_tile.json 2> $O/j13_bench_tile.err; echo "rc=$?" >> $O/j13_bench_tile.err
LUXB200_KS_TILE=0 timeout 600 python bench.py --steps 3 --warmup 3 --no-extras --no-strong --no-cpu-baseline > $O/j13_bench_gather.json 2> $O/j13_bench_gather.err; echo "rc=$?" >> $O/j13_bench_gather.err
python - <<'PY'
import json
for f in ("j13_bench_tile","j13_bench_gather"):
    try:
        d=json.loads(open("gpurun_out/%s.json"%f).read().strip().splitlines()[-1])
        print(f, d["value"], d["e2e"]["value"], d["ks_ms_per_step"], d["br_ms_per_step"], d["decrypt_errors"])
    except Exception as e:
        print(f, "failed", e)
PY
timeout 900 ncu --set full --clock-control none --import-source on -k regex:ks_ -c 2 -o $O/j13_ks_tile python bench.py --steps 1 --warmup 1 --no-extras --no-strong --no-cpu-baseline > $O/j13_ncu.log 2>&1; echo "ncu rc=$?" >> $O/j13_ncu.log
tail -2 $O/j13_ncu.log
