#!/bin/bash
# r2 job 18: full bench line of the current build
O=gpurun_out; mkdir -p $O
timeout 900 python bench.py --steps 3 --warmup 3 > $O/j18_bench.json 2> $O/j18_bench.err; echo "rc=$?" >> $O/j18_bench.err
tail -n 2 $O/j18_bench.err
python - <<'PY'
import json
d=json.loads(open("gpurun_out/j18_bench.json").read().strip().splitlines()[-1])
print(d["value"], d["e2e"]["value"], d["ks_ms_per_step"], d["br_ms_per_step"], d["decrypt_errors"])
for k in ("lmkcdey","ap"):
    print(k, d[k]["value"], d[k]["e2e"]["value"], d[k]["roofline"]["frac"], d[k]["decrypt_errors"])
print(d["ntt_n65536"]["forward"], d["ntt_n65536"]["inverse"])
PY
