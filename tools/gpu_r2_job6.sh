#!/bin/bash
# A/B: scheduled-kernel variants (LMKCDEY / AP), 14208 gates = whole waves for 6 and 8 ciphertexts per CTA; + pending tests
O=gpurun_out
for v in main cts8 cts8pair cts6pair; do
  for m in LMKCDEY AP; do
    if [ $v = main ]; then so=fhe_b200/libluxb200.so; else so=fhe_b200/_dev/libluxb200_$v.so; fi
    LUXB200_SO=$so python tools/prof_sched.py $m 14208 > $O/ab2_${m}_$v.log 2>&1; echo "$v $m: $(tail -2 $O/ab2_${m}_$v.log | head -1 | cut -c1-200)"
  done
done
( time python -m pytest tests/test_circuit.py tests/test_host_cpp.py -m gpu -q -x -k "evalfunc or largeprec" ) > $O/r2_t6.log 2>&1; tail -8 $O/r2_t6.log
