#!/bin/bash
# final ncu --set full capture of the three frame kernels: one frame on all SMs, two at the bench's 9-SM share
mkdir -p gpurun_out
python tools/ncu_frame.py > gpurun_out/r2s39_ncu_plain.log 2>&1 && timeout 1200 ncu --set full --clock-control none --import-source on -k "regex:vote_frame_kernel|fused_loop_kernel|frame_prefilter_kernel" -s 6 -c 9 -o gpurun_out/r2s39_frame python tools/ncu_frame.py > gpurun_out/r2s39_ncu.log 2>&1; echo "ncu rc=$?"
cat gpurun_out/r2s39_ncu_plain.log
