#!/bin/bash
# round-2 GPU session 2: the frame fast path (cluster K0 + device-side ctor + one wait)
mkdir -p gpurun_out
export CUDA_DEVICE_MAX_CONNECTIONS=32
timeout 900 python -m pytest tests/test_fast_frames_gpu.py tests/test_pipeline_gpu.py tests/test_stream_vote_gpu.py -q > gpurun_out/r2s2_fast.log 2>&1; echo "fast rc=$?" >> gpurun_out/r2s2_fast.log
timeout 900 python -m pytest tests/test_replay_parity_gpu.py tests/test_segments_gpu.py tests/test_world_gpu.py tests/test_cloud2_gpu.py tests/test_fused_paths_gpu.py -x -q > gpurun_out/r2s2_parity.log 2>&1; echo "parity rc=$?" >> gpurun_out/r2s2_parity.log
timeout 600 python bench.py --steps 10 --warmup 3 --no-direction-sharded > gpurun_out/r2s2_bench.json 2> gpurun_out/r2s2_bench.err; echo "bench rc=$?" >> gpurun_out/r2s2_bench.err
for sh in 0 18; do for fr in 0 7 15; do timeout 120 python tools/prof_fused.py $sh $fr > gpurun_out/r2s2_prof_share${sh}_frame${fr}.txt 2>&1; done; done
tail -n 4 gpurun_out/r2s2_fast.log gpurun_out/r2s2_parity.log gpurun_out/r2s2_bench.err
timeout 600 python -m pytest tests/test_hough_gpu.py tests/test_loop_gpu.py tests/test_fullsize_gpu.py tests/test_golden_gpu.py -x -q > gpurun_out/r2s2_rest.log 2>&1; echo "rest rc=$?" >> gpurun_out/r2s2_rest.log
timeout 600 python tools/stream_ab.py > gpurun_out/r2s2_stream_ab.txt 2>&1
tail -n 4 gpurun_out/r2s2_rest.log; cat gpurun_out/r2s2_stream_ab.txt
