#!/bin/bash
# r2 job 40: hybrid key switching composite on the device vs the oracle (N up to 2^16); all RNS tests
O=gpurun_out; mkdir -p $O
timeout 1200 python -m pytest tests/test_gpu_parity.py tests/test_zz_no_alloc.py -m gpu -q -k "hybrid or rns or static" --durations=5 > $O/j40_pytest.log 2>&1; echo "pytest rc=$?" >> $O/j40_pytest.log
tail -n 25 $O/j40_pytest.log | cut -c1-250
