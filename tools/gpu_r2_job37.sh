#!/bin/bash
# r2 job 37: gate circuit on STD256 (N = 2048) against the oracle on every wire
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests/test_circuit.py -m gpu -q -k "STD256" --durations=3 > $O/j37_pytest.log 2>&1; echo "pytest rc=$?" >> $O/j37_pytest.log
tail -n 8 $O/j37_pytest.log | cut -c1-250
