#!/bin/bash
mkdir -p gpurun_out
T=${1:-r2s24}
A="--no-cpu-baseline --no-direction-sharded"
python bench.py $A > gpurun_out/${T}_n1.json 2> gpurun_out/${T}_n1.err; echo "n1 rc=$?"
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 $A > gpurun_out/${T}_n2.json 2> gpurun_out/${T}_n2.err; echo "n2 rc=$?"
