#!/bin/bash
# bucketed un-vote: parity suites first, then the bench and the share-9 phase tables
mkdir -p gpurun_out
T=${1:-r2s11}
timeout 1500 python -m pytest tests/test_replay_parity_gpu.py tests/test_fast_frames_gpu.py tests/test_fused_paths_gpu.py tests/test_loop_gpu.py tests/test_pipeline_gpu.py -x -q -m gpu > gpurun_out/${T}_tests.log 2>&1; echo "rc=$?" >> gpurun_out/${T}_tests.log
timeout 900 python bench.py --no-direction-sharded > gpurun_out/${T}_bench.json 2> gpurun_out/${T}_bench.err; echo "bench rc=$?" >> gpurun_out/${T}_bench.err
for f in 0 7 15; do timeout 300 python tools/prof_fused.py 9 $f > gpurun_out/${T}_loop_share9_frame$f.txt 2>&1; done
timeout 300 python tools/prof_fused.py 0 7 > gpurun_out/${T}_loop_share0_frame7.txt 2>&1
tail -n 3 gpurun_out/${T}_tests.log gpurun_out/${T}_bench.err; tail -n 1 gpurun_out/${T}_loop_share9_frame*.txt gpurun_out/${T}_loop_share0_frame7.txt
