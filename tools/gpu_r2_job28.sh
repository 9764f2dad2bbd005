#!/bin/bash
# r2 job 28: final build — the whole -m gpu suite twice (flakiness check), smoke(), the bench line
O=gpurun_out; mkdir -p $O
for i in 1 2; do
  timeout 3000 python -m pytest tests -m gpu -q -x --durations=8 > $O/j28_pytest_full_$i.log 2>&1; echo "full rc=$?" >> $O/j28_pytest_full_$i.log
  tail -n 3 $O/j28_pytest_full_$i.log | cut -c1-250
done
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/j28_smoke.log 2>&1; echo "smoke rc=$?" >> $O/j28_smoke.log; tail -n 2 $O/j28_smoke.log
timeout 900 python bench.py --steps 3 --warmup 3 > $O/j28_bench.json 2> $O/j28_bench.err; echo "rc=$?" >> $O/j28_bench.err
tail -n 2 $O/j28_bench.err
python - <<'PY'
import json
d=json.loads(open("gpurun_out/j28_bench.json").read().strip().splitlines()[-1])
print(d["value"], d["e2e"]["value"], d["ks_ms_per_step"], d["br_ms_per_step"], d["decrypt_errors"], d["roofline"]["frac"])
for k,v in d["n2048"].items():
    print(k, v["value"], v.get("roofline",{}).get("frac"))
for k in ("lmkcdey","ap"):
    print(k, d[k]["value"], d[k]["e2e"]["value"], d[k]["roofline"]["frac"], d[k]["decrypt_errors"])
PY
