#!/bin/bash
# the allocation-free entry point: alone in a fresh process, then at the end of the whole suite; then the
# final concurrency sweep and the ncu launch list of the default bench command
mkdir -p gpurun_out
T=r2s9
timeout 600 python -m pytest tests/test_zz_no_alloc.py -x -q -m gpu > gpurun_out/${T}_noalloc_alone.log 2>&1; echo "rc=$?" >> gpurun_out/${T}_noalloc_alone.log
timeout 2400 python -m pytest tests -x -q -m gpu > gpurun_out/${T}_tests.log 2>&1; echo "rc=$?" >> gpurun_out/${T}_tests.log
timeout 900 python tools/concurrency_sweep.py 6 12:18,16:9,16:12,24:9,24:6,20:7,32:4 > gpurun_out/${T}_sweep.txt 2>&1
timeout 600 python bench.py --steps 2 --warmup 1 --no-direction-sharded > gpurun_out/${T}_bench_short.json 2> gpurun_out/${T}_bench_short.err \
  && timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/${T}_launches.csv \
     python bench.py --steps 2 --warmup 1 --no-direction-sharded > gpurun_out/${T}_ncu.log 2>&1
tail -n 3 gpurun_out/${T}_noalloc_alone.log gpurun_out/${T}_tests.log gpurun_out/${T}_sweep.txt
