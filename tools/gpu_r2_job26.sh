#!/bin/bash
# r2 job 26: the N = 2048 sets against the LIVE reference through the C++ host layer (host_parity)
O=gpurun_out; mkdir -p $O
timeout 3000 python -m pytest tests/test_host_cpp.py -m gpu -q -k "STD256 or STD192" --durations=5 > $O/j26_pytest.log 2>&1; echo "pytest rc=$?" >> $O/j26_pytest.log
tail -n 30 $O/j26_pytest.log | cut -c1-300
