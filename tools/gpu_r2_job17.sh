#!/bin/bash
# r2 job 17: u32 conditional subtraction as min() in br_wide.cuh (parity of the wide sets, keygen), STD256 LMKCDEY back on the
# shared-memory kernel, throughput of the wide sets, full bench line (pinned e2e for the LMKCDEY / AP extras)
O=gpurun_out; mkdir -p $O
timeout 1500 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "wide or keygen or largeprec" --durations=5 > $O/j17_pytest.log 2>&1; echo "pytest rc=$?" >> $O/j17_pytest.log
tail -n 9 $O/j17_pytest.log
timeout 900 python tools/bench_wide.py 1184 > $O/j17_wide_bench.jsonl 2> $O/j17_wide_bench.err; echo "rc=$?" >> $O/j17_wide_bench.err
cut -c1-220 $O/j17_wide_bench.jsonl; tail -n 2 $O/j17_wide_bench.err
timeout 900 python bench.py --steps 3 --warmup 3 > $O/j17_bench.json 2> $O/j17_bench.err; echo "rc=$?" >> $O/j17_bench.err
tail -n 2 $O/j17_bench.err
python - <<'PY'
import json
d=json.loads(open("gpurun_out/j17_bench.json").read().strip().splitlines()[-1])
print(d["value"], d["e2e"]["value"], d["ks_ms_per_step"], d["br_ms_per_step"], d["decrypt_errors"])
for k in ("lmkcdey","ap"):
    print(k, d[k]["value"], d[k]["e2e"]["value"], d[k]["roofline"]["frac"], d[k]["decrypt_errors"])
PY
