#!/bin/bash
# r2 job 14: ks_tile_kernel A/B (16 warps x 16 ciphertexts vs 32 warps x 8), parity of both, ncu of the winner; ncu of brw_kernel
O=gpurun_out; mkdir -p $O
for cpw in 16 8; do
  LUXB200_KS_CPW=$cpw timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "keyswitch_tile" > $O/j14_pytest_cpw$cpw.log 2>&1; echo "pytest rc=$?" >> $O/j14_pytest_cpw$cpw.log; tail -2 $O/j14_pytest_cpw$cpw.log
  LUXB200_KS_CPW=$cpw timeout 600 python bench.py --steps 3 --warmup 3 --no-extras --no-strong --no-cpu-baseline --no-ntt > $O/j14_bench_cpw$cpw.json 2> $O/j14_bench_cpw$cpw.err; echo "rc=$?" >> $O/j14_bench_cpw$cpw.err
done
python - <<'PY'
import json
for f in ("j14_bench_cpw16","j14_bench_cpw8"):
    try:
        d=json.loads(open("gpurun_out/%s.json"%f).read().strip().splitlines()[-1])
        print(f, d["value"], d["e2e"]["value"], d["ks_ms_per_step"], d["br_ms_per_step"], d["decrypt_errors"])
    except Exception as e:
        print(f, "failed", e)
PY
LUXB200_KS_CPW=8 timeout 900 ncu --set full --clock-control none --import-source on -k regex:ks_tile -c 1 -o $O/j14_ks_tile_cpw8 python bench.py --steps 1 --warmup 1 --no-extras --no-strong --no-cpu-baseline --no-ntt > $O/j14_ncu_ks.log 2>&1; echo "ncu rc=$?" >> $O/j14_ncu_ks.log
timeout 600 python tools/prof_wide.py STD256 444 > $O/j14_wide_plain.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:brw_kernel -c 1 -o $O/j14_brw_std256 python tools/prof_wide.py STD256 444 > $O/j14_ncu_brw.log 2>&1; echo "ncu rc=$?" >> $O/j14_ncu_brw.log
tail -2 $O/j14_wide_plain.log $O/j14_ncu_ks.log $O/j14_ncu_brw.log
