#!/bin/bash
# K0 phase stamps and the loop's phase table at the bench's SM share (9)
mkdir -p gpurun_out
T=r2s10
timeout 300 python tools/k0_phases.py 0 > gpurun_out/${T}_k0_share0.txt 2>&1
timeout 300 python tools/k0_phases.py 9 > gpurun_out/${T}_k0_share9.txt 2>&1
for f in 0 7 15; do timeout 300 python tools/prof_fused.py 9 $f > gpurun_out/${T}_loop_share9_frame$f.txt 2>&1; done
tail -n 4 gpurun_out/${T}_k0_share0.txt gpurun_out/${T}_k0_share9.txt; tail -n 2 gpurun_out/${T}_loop_share9_frame7.txt
