#!/bin/bash
# r2 job 12 (gpurun --gpus 8): the 8-GPU bench line incl. the strong-scaling block (65 536 gates in total), and the reference-side
# call lux::fhe::EvalBinGateBatch through ONE BackendB200 over 8 devices
O=gpurun_out; mkdir -p $O
nvidia-smi --query-gpu=index,name,clocks.sm --format=csv > $O/j12_smi.txt
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 8 --steps 3 --warmup 3 --no-extras > $O/j12_bench8.json 2> $O/j12_bench8.err; echo "bench8 rc=$?" >> $O/j12_bench8.err
export LD_LIBRARY_PATH=$PWD/fhe_b200:$PWD/oracle/_ref:$PWD/host/_build:$LD_LIBRARY_PATH
timeout 600 host/_build/host_parity STD128 9 2 --devices 8 --e2e 131072 > $O/j12_e2e_cpp_8dev.txt 2>&1; echo "parity8 rc=$?" >> $O/j12_e2e_cpp_8dev.txt
timeout 300 host/_build/host_parity STD128 9 0 --devices 8 --e2e 65536 > $O/j12_e2e_cpp_8dev_strong.txt 2>&1; echo "parity8s rc=$?" >> $O/j12_e2e_cpp_8dev_strong.txt
tail -n 3 $O/j12_bench8.err $O/j12_e2e_cpp_8dev.txt $O/j12_e2e_cpp_8dev_strong.txt
