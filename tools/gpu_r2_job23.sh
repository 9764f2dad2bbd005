#!/bin/bash
# r2 job 23: full -m gpu suite with the quad kernel build, then the bench line and the N = 2048 throughput table
O=gpurun_out; mkdir -p $O
timeout 3000 python -m pytest tests -m gpu -q --durations=10 > $O/j23_pytest_full.log 2>&1; echo "full rc=$?" >> $O/j23_pytest_full.log
tail -n 22 $O/j23_pytest_full.log | cut -c1-250
timeout 900 python bench.py --steps 3 --warmup 3 > $O/j23_bench.json 2> $O/j23_bench.err; echo "rc=$?" >> $O/j23_bench.err
tail -n 2 $O/j23_bench.err
python - <<'PY'
import json
d=json.loads(open("gpurun_out/j23_bench.json").read().strip().splitlines()[-1])
print(d["value"], d["e2e"]["value"], d["ks_ms_per_step"], d["br_ms_per_step"], d["decrypt_errors"])
for k,v in d["n2048"].items():
    print(k, v["value"], v.get("roofline",{}).get("frac"))
for k in ("lmkcdey","ap"):
    print(k, d[k]["value"], d[k]["e2e"]["value"], d[k]["roofline"]["frac"], d[k]["decrypt_errors"])
PY
timeout 900 python tools/bench_wide.py 1776 > $O/j23_wide_bench.jsonl 2> $O/j23_wide_bench.err; echo "rc=$?" >> $O/j23_wide_bench.err
cut -c1-220 $O/j23_wide_bench.jsonl
