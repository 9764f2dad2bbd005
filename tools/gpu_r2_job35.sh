#!/bin/bash
# r2 job 35: launch list + DRAM bytes of the final build's bench step (ncu pass; nothing printed under ncu is a bench value)
O=gpurun_out; mkdir -p $O
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-extras --no-strong --no-ntt"
timeout 600 $CMD > $O/j35_plain.json 2> $O/j35_plain.err; echo "plain rc=$?" >> $O/j35_plain.err; tail -n 1 $O/j35_plain.err
timeout 1200 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 400 --csv --log-file $O/j35_launches.csv $CMD > $O/j35_ncu.log 2>&1; echo "ncu rc=$?" >> $O/j35_ncu.log; tail -n 1 $O/j35_ncu.log
wc -l $O/j35_launches.csv
