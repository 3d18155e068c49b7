#!/bin/bash
# split-phase replay workers: parity, then the bench with one thread per pipeline and with 4 / 8 threads
mkdir -p gpurun_out
T=${1:-r2s17}
timeout 1500 python -m pytest tests/test_replay_parity_gpu.py tests/test_world_gpu.py tests/test_fast_frames_gpu.py tests/test_pipeline_gpu.py -x -q -m gpu > gpurun_out/${T}_tests.log 2>&1; echo "rc=$?" >> gpurun_out/${T}_tests.log
for th in -1 8 4 2; do
  timeout 600 python bench.py --no-direction-sharded --no-cpu-baseline --replay-threads $th > gpurun_out/${T}_bench_t$th.json 2> gpurun_out/${T}_bench_t$th.err; echo "bench rc=$?" >> gpurun_out/${T}_bench_t$th.err
done
tail -n 3 gpurun_out/${T}_tests.log; tail -n 1 gpurun_out/${T}_bench_t*.err
