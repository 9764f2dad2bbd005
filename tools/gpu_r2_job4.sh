#!/bin/bash
# 2-GPU box: multi-device BackendB200 (one process, keys peer-to-peer) + the torchrun bench at N = 2
O=gpurun_out
nvidia-smi --query-gpu=index,name --format=csv,noheader > $O/r2_2gpu.txt
( time python -m pytest tests/test_host_cpp.py -m gpu -q -x -k "multi_device" ) > $O/r2_t4.log 2>&1; tail -12 $O/r2_t4.log
host/_build/host_parity STD128 6 0 --devices 2 --e2e 131072 > $O/r2_e2e_cpp_2gpu.log 2>&1; grep -E "e2e_cpp|backend:" $O/r2_e2e_cpp_2gpu.log
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 2 --steps 3 --warmup 3 > $O/r2_bench_2gpu.log 2>&1; tail -1 $O/r2_bench_2gpu.log | cut -c1-400
