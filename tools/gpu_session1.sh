#!/bin/bash
# round-2 GPU session 1: parity at bench concurrency, full suite, bench, sanitizers
mkdir -p gpurun_out
export CUDA_DEVICE_MAX_CONNECTIONS=32
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > gpurun_out/r2s1_smi.txt 2>&1
timeout 900 python -m pytest tests/test_replay_parity_gpu.py -x -q > gpurun_out/r2s1_parity.log 2>&1; echo "parity rc=$?" >> gpurun_out/r2s1_parity.log
timeout 1500 python -m pytest tests -x -q -m gpu --deselect tests/test_replay_parity_gpu.py > gpurun_out/r2s1_tests.log 2>&1; echo "tests rc=$?" >> gpurun_out/r2s1_tests.log
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/r2s1_bench.json 2> gpurun_out/r2s1_bench.err; echo "bench rc=$?" >> gpurun_out/r2s1_bench.err
timeout 900 compute-sanitizer --tool racecheck --racecheck-report all python tools/sanitize_frame.py > gpurun_out/r2s1_racecheck.log 2>&1; echo "rc=$?" >> gpurun_out/r2s1_racecheck.log
timeout 900 compute-sanitizer --tool memcheck python tools/sanitize_frame.py > gpurun_out/r2s1_memcheck.log 2>&1; echo "rc=$?" >> gpurun_out/r2s1_memcheck.log
timeout 600 compute-sanitizer --tool synccheck python tools/sanitize_frame.py > gpurun_out/r2s1_synccheck.log 2>&1; echo "rc=$?" >> gpurun_out/r2s1_synccheck.log
H3D_SAN_GENERAL=1 timeout 900 compute-sanitizer --tool racecheck --racecheck-report all python tools/sanitize_frame.py > gpurun_out/r2s1_racecheck_general.log 2>&1; echo "rc=$?" >> gpurun_out/r2s1_racecheck_general.log
tail -3 gpurun_out/r2s1_parity.log gpurun_out/r2s1_tests.log gpurun_out/r2s1_bench.err gpurun_out/r2s1_racecheck.log gpurun_out/r2s1_memcheck.log gpurun_out/r2s1_synccheck.log gpurun_out/r2s1_racecheck_general.log
