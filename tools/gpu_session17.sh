#!/bin/bash
mkdir -p gpurun_out
T=${1:-r2s27}
timeout 1500 python -m pytest tests/test_fullsize_gpu.py tests/test_loop_gpu.py tests/test_pointcloud_gpu.py tests/test_golden_gpu.py tests/test_sharding_gpu.py -x -q -m gpu > gpurun_out/${T}_tests.log 2>&1; echo "rc=$?" >> gpurun_out/${T}_tests.log
timeout 900 python bench.py --no-cpu-baseline > gpurun_out/${T}_bench.json 2> gpurun_out/${T}_bench.err; echo "bench rc=$?" >> gpurun_out/${T}_bench.err
python tools/c4_once.py > gpurun_out/${T}_plain.log 2>&1 && timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file gpurun_out/${T}_c4_launches.csv python tools/c4_once.py > gpurun_out/${T}_ncu.log 2>&1
tail -n 3 gpurun_out/${T}_tests.log gpurun_out/${T}_bench.err
