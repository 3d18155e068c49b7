#!/bin/bash
# what costs 13 % per GPU at N = 2?  the same torchrun launch with gloo plumbing, lazy NCCL, eager NCCL, NVLS off
mkdir -p gpurun_out
T=r2s22
run() { # name, env...
  name=$1; shift
  env "$@" python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --no-cpu-baseline --no-direction-sharded > gpurun_out/${T}_$name.json 2> gpurun_out/${T}_$name.err
  echo "$name rc=$?"
}
run gloo H3D_BENCH_BACKEND=gloo
run lazy H3D_BENCH_BACKEND=nccl-lazy
run eager H3D_BENCH_BACKEND=nccl
run nvls0 H3D_BENCH_BACKEND=nccl NCCL_NVLS_ENABLE=0
run p2p0 H3D_BENCH_BACKEND=nccl NCCL_P2P_DISABLE=1 NCCL_SHM_DISABLE=0
