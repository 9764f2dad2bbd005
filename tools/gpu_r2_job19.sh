#!/bin/bash
# r2 job 19: one-wave first chunk in the pipelined host path; N = 2048 block in the bench extras
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "full_batch or keyswitch_tile or golden" --durations=3 > $O/j19_pytest.log 2>&1; echo "pytest rc=$?" >> $O/j19_pytest.log
tail -n 6 $O/j19_pytest.log
timeout 900 python bench.py --steps 3 --warmup 3 > $O/j19_bench.json 2> $O/j19_bench.err; echo "rc=$?" >> $O/j19_bench.err
tail -n 2 $O/j19_bench.err
python - <<'PY'
import json
d=json.loads(open("gpurun_out/j19_bench.json").read().strip().splitlines()[-1])
print(d["value"], d["e2e"]["value"], d["ks_ms_per_step"], d["br_ms_per_step"], d["decrypt_errors"])
for k,v in d["n2048"].items():
    print(k, v["value"], v.get("roofline",{}).get("frac"))
PY
