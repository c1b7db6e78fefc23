#!/bin/bash
# Two-GPU verification: the contract bench line at N=2 (independent batches), the row-sharded C5 workload, the shard tests.
mkdir -p gpurun_out
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 30 --warmup 3 > gpurun_out/n2_bench.log 2>&1; echo "rc=$?" >> gpurun_out/n2_bench.log
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --workload c5 --steps 10 --warmup 3 > gpurun_out/n2_c5.log 2>&1; echo "rc=$?" >> gpurun_out/n2_c5.log
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 bench.py --impl reference --gpus 2 --steps 2 --warmup 1 > gpurun_out/n2_ref.log 2>&1; echo "rc=$?" >> gpurun_out/n2_ref.log
timeout 600 python -m pytest tests/test_gpu_shard.py -q > gpurun_out/n2_shard.log 2>&1; echo "rc=$?" >> gpurun_out/n2_shard.log
grep '^{\|rc=' gpurun_out/n2_bench.log | cut -c1-400; grep '^{\|rc=' gpurun_out/n2_c5.log | cut -c1-400; grep '^{\|rc=' gpurun_out/n2_ref.log | cut -c1-300; tail -3 gpurun_out/n2_shard.log
