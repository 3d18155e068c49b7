#!/bin/bash
mkdir -p gpurun_out
export CUDA_DEVICE_MAX_CONNECTIONS=32
timeout 900 python -m pytest tests/test_fast_frames_gpu.py tests/test_pipeline_gpu.py tests/test_replay_parity_gpu.py tests/test_fused_paths_gpu.py tests/test_segments_gpu.py tests/test_stream_vote_gpu.py tests/test_world_gpu.py -x -q > gpurun_out/r2s4_tests.log 2>&1; echo "tests rc=$?" >> gpurun_out/r2s4_tests.log
timeout 300 python tools/track_ab.py > gpurun_out/r2s4_track_ab.txt 2>&1
timeout 600 python tools/stream_ab.py > gpurun_out/r2s4_stream_ab.txt 2>&1
for sh in 0 18; do for fr in 7; do timeout 120 python tools/prof_fused.py $sh $fr > gpurun_out/r2s4_prof_share${sh}_frame${fr}.txt 2>&1; done; done
export CUDA_DEVICE_MAX_CONNECTIONS=64
timeout 900 python tools/concurrency_sweep.py 6 12:18,16:9,16:12,24:6,24:9,32:4,32:6,40:4,48:3 > gpurun_out/r2s4_sweep.txt 2>&1
export CUDA_DEVICE_MAX_CONNECTIONS=32
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/r2s4_bench.json 2> gpurun_out/r2s4_bench.err; echo "bench rc=$?" >> gpurun_out/r2s4_bench.err
timeout 900 python -m pytest tests -x -q -m gpu --deselect tests/test_fast_frames_gpu.py --deselect tests/test_pipeline_gpu.py --deselect tests/test_replay_parity_gpu.py --deselect tests/test_fused_paths_gpu.py --deselect tests/test_segments_gpu.py --deselect tests/test_stream_vote_gpu.py --deselect tests/test_world_gpu.py > gpurun_out/r2s4_rest.log 2>&1; echo "rest rc=$?" >> gpurun_out/r2s4_rest.log
tail -n 3 gpurun_out/r2s4_tests.log gpurun_out/r2s4_bench.err gpurun_out/r2s4_rest.log; cat gpurun_out/r2s4_track_ab.txt gpurun_out/r2s4_stream_ab.txt gpurun_out/r2s4_sweep.txt
