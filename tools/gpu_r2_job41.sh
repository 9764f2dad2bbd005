#!/bin/bash
# r2 job 41: hybrid key switch tests after the _dev refactor; full bench line with the hks_n65536 block
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -k "hybrid or rns" > $O/j41_pytest.log 2>&1; echo "pytest rc=$?" >> $O/j41_pytest.log; tail -n 3 $O/j41_pytest.log | cut -c1-200
( time timeout 900 python bench.py --steps 3 --warmup 3 > $O/j41_bench.json 2> $O/j41_bench.err ) 2> $O/j41_time.txt; echo "rc=$?" >> $O/j41_bench.err
tail -n 2 $O/j41_bench.err; cat $O/j41_time.txt
python - <<'PY'
import json
d=json.loads(open("gpurun_out/j41_bench.json").read().strip().splitlines()[-1])
print(d["value"], d["e2e"]["value"], d["decrypt_errors"])
print(json.dumps(d["hks_n65536"])[:900])
for k,v in d["n2048"].items():
    print(k, v["value"], v.get("roofline",{}).get("frac"))
PY
