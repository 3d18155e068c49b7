#!/bin/bash
# source-level ncu capture of the frame vote (TRACK form, bench frame 7: 2848 points, 6 slots), all SMs and an 18-SM share
mkdir -p gpurun_out
python tools/ncu_frame.py 7 > gpurun_out/r2s14_plain.log 2>&1 && timeout 900 ncu --set full --clock-control none --import-source on -k "regex:vote_frame_kernel" -c 5 -o gpurun_out/r2s14_vote python tools/ncu_frame.py 7 > gpurun_out/r2s14_ncu.log 2>&1; echo "ncu rc=$?" >> gpurun_out/r2s14_ncu.log
cat gpurun_out/r2s14_plain.log; tail -3 gpurun_out/r2s14_ncu.log
