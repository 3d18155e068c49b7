#!/bin/bash
mkdir -p gpurun_out
export CUDA_DEVICE_MAX_CONNECTIONS=32
timeout 900 python -m pytest tests/test_fast_frames_gpu.py tests/test_pipeline_gpu.py tests/test_replay_parity_gpu.py tests/test_fused_paths_gpu.py tests/test_segments_gpu.py -x -q > gpurun_out/r2s7_tests.log 2>&1; echo "tests rc=$?" >> gpurun_out/r2s7_tests.log
timeout 300 python tools/track_ab.py > gpurun_out/r2s7_track_ab.txt 2>&1
timeout 120 python tools/prof_fused.py 18 7 > gpurun_out/r2s7_prof_share18_frame7.txt 2>&1
timeout 120 python tools/prof_fused.py 0 7 > gpurun_out/r2s7_prof_share0_frame7.txt 2>&1
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-direction-sharded > gpurun_out/r2s7_bench.json 2> gpurun_out/r2s7_bench.err; echo "bench rc=$?" >> gpurun_out/r2s7_bench.err
export CUDA_DEVICE_MAX_CONNECTIONS=64
timeout 600 python tools/concurrency_sweep.py 6 12:18,16:9,16:12,24:9,20:14,12:24,10:29,8:37 > gpurun_out/r2s7_sweep.txt 2>&1
tail -n 3 gpurun_out/r2s7_tests.log gpurun_out/r2s7_bench.err; cut -c1-330 gpurun_out/r2s7_track_ab.txt; cat gpurun_out/r2s7_sweep.txt
