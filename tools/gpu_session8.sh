#!/bin/bash
mkdir -p gpurun_out
python tools/ncu_vote.py > gpurun_out/r2s8_plain.log 2>&1 && timeout 900 ncu --set full --clock-control none --import-source on -k "regex:vote_frame_kernel|vote_resident_kernel" -c 6 -o gpurun_out/r2s8_vote python tools/ncu_vote.py > gpurun_out/r2s8_ncu.log 2>&1; echo "ncu rc=$?" >> gpurun_out/r2s8_ncu.log
cat gpurun_out/r2s8_plain.log; tail -3 gpurun_out/r2s8_ncu.log
