#!/bin/bash
# r2 job 43: final build — the whole -m gpu suite (-x, like the driver), smoke()
O=gpurun_out; mkdir -p $O
timeout 3000 python -m pytest tests -m gpu -q -x --durations=8 > $O/j43_pytest_full.log 2>&1; echo "full rc=$?" >> $O/j43_pytest_full.log
tail -n 12 $O/j43_pytest_full.log | cut -c1-250
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/j43_smoke.log 2>&1; echo "smoke rc=$?" >> $O/j43_smoke.log; tail -n 2 $O/j43_smoke.log
