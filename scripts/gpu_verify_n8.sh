#!/bin/bash
# Eight-GPU verification: the contract bench line (independent batches) and the row-sharded C5 workload.
mkdir -p gpurun_out
N=${1:-8}
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus $N --steps 30 --warmup 3 > gpurun_out/n${N}_bench.log 2>&1; echo "rc=$?" >> gpurun_out/n${N}_bench.log
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29522 bench.py --gpus $N --workload c5 --steps 10 --warmup 3 > gpurun_out/n${N}_c5.log 2>&1; echo "rc=$?" >> gpurun_out/n${N}_c5.log
grep '^{\|rc=' gpurun_out/n${N}_bench.log | cut -c1-260; grep '^{\|rc=' gpurun_out/n${N}_c5.log | cut -c1-260
