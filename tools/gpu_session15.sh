#!/bin/bash
mkdir -p gpurun_out
T=r2s23
A="--no-cpu-baseline --no-direction-sharded"
OMP_NUM_THREADS=1 python bench.py $A > gpurun_out/${T}_omp1.json 2> gpurun_out/${T}_omp1.err; echo "omp1 rc=$?"
python -m torch.distributed.run --nnodes=1 --nproc-per-node 1 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 1 $A > gpurun_out/${T}_torchrun1.json 2> gpurun_out/${T}_torchrun1.err; echo "torchrun1 rc=$?"
RANK=0 WORLD_SIZE=1 LOCAL_RANK=1 python bench.py $A > gpurun_out/${T}_gpu1.json 2> gpurun_out/${T}_gpu1.err; echo "gpu1 rc=$?"
