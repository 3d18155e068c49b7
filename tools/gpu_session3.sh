#!/bin/bash
# round-2 GPU session 3: fast path v2 (device-side P, batched K0, TRACK vote, warp-per-record un-vote, inlier ILP)
mkdir -p gpurun_out
export CUDA_DEVICE_MAX_CONNECTIONS=32
timeout 900 python -m pytest tests/test_fast_frames_gpu.py tests/test_pipeline_gpu.py tests/test_replay_parity_gpu.py tests/test_fused_paths_gpu.py tests/test_segments_gpu.py -x -q > gpurun_out/r2s3_tests.log 2>&1; echo "tests rc=$?" >> gpurun_out/r2s3_tests.log
timeout 600 python bench.py --steps 10 --warmup 3 --no-direction-sharded --no-cpu-baseline > gpurun_out/r2s3_bench.json 2> gpurun_out/r2s3_bench.err; echo "bench rc=$?" >> gpurun_out/r2s3_bench.err
timeout 300 python tools/track_ab.py > gpurun_out/r2s3_track_ab.txt 2>&1
for sh in 0 18; do for fr in 0 7; do timeout 120 python tools/prof_fused.py $sh $fr > gpurun_out/r2s3_prof_share${sh}_frame${fr}.txt 2>&1; done; done
timeout 900 python tools/concurrency_sweep.py 6 12:18,16:18,16:12,24:12,24:9,32:9,12:24,20:14,16:9,24:6 > gpurun_out/r2s3_sweep.txt 2>&1
timeout 600 python -m pytest tests/test_fullsize_gpu.py tests/test_sharding_gpu.py tests/test_loop_gpu.py -x -q > gpurun_out/r2s3_big.log 2>&1; echo "big rc=$?" >> gpurun_out/r2s3_big.log
timeout 300 python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/r2s3_bench_sharded.json 2> gpurun_out/r2s3_bench_sharded.err
tail -n 3 gpurun_out/r2s3_tests.log gpurun_out/r2s3_bench.err gpurun_out/r2s3_big.log; cat gpurun_out/r2s3_track_ab.txt; tail -20 gpurun_out/r2s3_sweep.txt
