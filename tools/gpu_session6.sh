#!/bin/bash
mkdir -p gpurun_out
export CUDA_DEVICE_MAX_CONNECTIONS=32
timeout 900 python -m pytest tests/test_fast_frames_gpu.py tests/test_pipeline_gpu.py tests/test_replay_parity_gpu.py tests/test_fused_paths_gpu.py tests/test_segments_gpu.py tests/test_dropin_gpu.py tests/test_cloud2_gpu.py tests/test_world_gpu.py -x -q > gpurun_out/r2s6_tests.log 2>&1; echo "tests rc=$?" >> gpurun_out/r2s6_tests.log
timeout 300 python tools/track_ab.py > gpurun_out/r2s6_track_ab.txt 2>&1
timeout 120 python tools/prof_fused.py 18 7 > gpurun_out/r2s6_prof_share18_frame7.txt 2>&1
timeout 120 python tools/prof_fused.py 0 7 > gpurun_out/r2s6_prof_share0_frame7.txt 2>&1
timeout 900 python bench.py --workload c3 > gpurun_out/r2s6_c3.json 2> gpurun_out/r2s6_c3.err; echo "c3 rc=$?" >> gpurun_out/r2s6_c3.err
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/r2s6_bench.json 2> gpurun_out/r2s6_bench.err; echo "bench rc=$?" >> gpurun_out/r2s6_bench.err
timeout 600 python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-direction-sharded > gpurun_out/r2s6_plain.log 2>&1 && timeout 1500 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/r2s6_bench_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-direction-sharded > gpurun_out/r2s6_ncu.log 2>&1; echo "ncu rc=$?" >> gpurun_out/r2s6_ncu.log
tail -n 3 gpurun_out/r2s6_tests.log gpurun_out/r2s6_bench.err gpurun_out/r2s6_c3.err gpurun_out/r2s6_ncu.log; cut -c1-330 gpurun_out/r2s6_track_ab.txt; cat gpurun_out/r2s6_c3.json | cut -c1-1500
