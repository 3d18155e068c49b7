#!/bin/bash
mkdir -p gpurun_out
export CUDA_DEVICE_MAX_CONNECTIONS=32
timeout 900 python -m pytest tests/test_fast_frames_gpu.py tests/test_pipeline_gpu.py tests/test_replay_parity_gpu.py tests/test_fused_paths_gpu.py tests/test_segments_gpu.py -x -q > gpurun_out/r2s5_tests.log 2>&1; echo "tests rc=$?" >> gpurun_out/r2s5_tests.log
timeout 300 python tools/track_ab.py > gpurun_out/r2s5_track_ab.txt 2>&1
timeout 120 python tools/prof_fused.py 18 7 > gpurun_out/r2s5_prof_share18_frame7.txt 2>&1
timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-direction-sharded > gpurun_out/r2s5_bench.json 2> gpurun_out/r2s5_bench.err; echo "bench rc=$?" >> gpurun_out/r2s5_bench.err
python tools/ncu_frame.py > gpurun_out/r2s5_ncu_plain.log 2>&1 && timeout 1200 ncu --set full --clock-control none --import-source on -k "regex:vote_frame_kernel|fused_loop_kernel|frame_prefilter_kernel" -s 6 -c 9 -o gpurun_out/r2s5_frame python tools/ncu_frame.py > gpurun_out/r2s5_ncu.log 2>&1; echo "ncu rc=$?" >> gpurun_out/r2s5_ncu.log
tail -n 3 gpurun_out/r2s5_tests.log gpurun_out/r2s5_bench.err gpurun_out/r2s5_ncu.log; cut -c1-300 gpurun_out/r2s5_track_ab.txt
