#!/bin/bash
# r2 job 32: final build — new pipelined N = 2048 test alone, then the whole -m gpu suite twice, smoke(), bench
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -k "n2048_pipelined" > $O/j32_new.log 2>&1; echo "rc=$?" >> $O/j32_new.log; tail -n 4 $O/j32_new.log | cut -c1-250
for i in 1 2; do
  timeout 3000 python -m pytest tests -m gpu -q -x --durations=6 > $O/j32_pytest_full_$i.log 2>&1; echo "full rc=$?" >> $O/j32_pytest_full_$i.log
  tail -n 3 $O/j32_pytest_full_$i.log | cut -c1-250
done
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/j32_smoke.log 2>&1; echo "smoke rc=$?" >> $O/j32_smoke.log; tail -n 2 $O/j32_smoke.log
timeout 900 python bench.py --steps 3 --warmup 3 > $O/j32_bench.json 2> $O/j32_bench.err; echo "rc=$?" >> $O/j32_bench.err
tail -n 2 $O/j32_bench.err
python - <<'PY'
import json
d=json.loads(open("gpurun_out/j32_bench.json").read().strip().splitlines()[-1])
print(d["value"], d["e2e"]["value"], d["ks_ms_per_step"], d["br_ms_per_step"], d["decrypt_errors"], d["roofline"]["frac"])
for k,v in d["n2048"].items():
    print(k, v["value"], v.get("roofline",{}).get("frac"))
PY
