#!/bin/bash
# r2 job 30 (gpurun --gpus 2): multi-device backend tests, the 2-GPU bench line, and the crash loop after the warm-up fix
O=gpurun_out; mkdir -p $O
timeout 1500 python -m pytest tests/test_host_cpp.py tests/test_gpu_parity.py -m gpu -q -k "multi_device or broadcast or two_gpu or 2gpu" --durations=5 > $O/j30_pytest.log 2>&1; echo "pytest rc=$?" >> $O/j30_pytest.log
tail -n 6 $O/j30_pytest.log | cut -c1-250
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29544 bench.py --gpus 2 --steps 3 --warmup 3 --no-extras > $O/j30_bench2.json 2> $O/j30_bench2.err; echo "bench2 rc=$?" >> $O/j30_bench2.err
tail -n 1 $O/j30_bench2.err
python - <<'PY'
import json
d=json.loads(open("gpurun_out/j30_bench2.json").read().strip().splitlines()[-1])
print(d["n_gpus"], d["value"], d["e2e"]["value"], d.get("strong_2p16",{}).get("value"), d.get("strong_2p16",{}).get("e2e"), d["decrypt_errors"])
PY
fails=0
for i in $(seq 1 60); do
  timeout 120 host/_build/host_parity TOY_LMKCDEY 8 1 --gpu-keygen > $O/j30_run.out 2> $O/j30_run.err; rc=$?
  if [ $rc -ne 0 ]; then fails=$((fails+1)); echo "=== run $i rc=$rc" >> $O/j30_crash.log; cat $O/j30_run.err >> $O/j30_crash.log; fi
done
echo "TOY_LMKCDEY --gpu-keygen after the fix: $fails failures of 60" | tee -a $O/j30_crash.log
timeout 600 host/_build/host_parity STD128 9 2 --devices 2 --e2e 131072 2>&1 | grep "e2e_cpp\"\|failed_checks" | cut -c1-200 | tee $O/j30_e2e_cpp_2dev.txt
