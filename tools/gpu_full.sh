#!/bin/bash
# full GPU suite + the bench line (what the driver runs at round end) + the CPU arm
mkdir -p gpurun_out
T=${1:-full}
timeout 2400 python -m pytest tests -x -q -m gpu > gpurun_out/${T}_tests.log 2>&1; echo "tests rc=$?" >> gpurun_out/${T}_tests.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${T}_smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/${T}_smoke.log
timeout 900 python bench.py > gpurun_out/${T}_bench.json 2> gpurun_out/${T}_bench.err; echo "bench rc=$?" >> gpurun_out/${T}_bench.err
timeout 900 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/${T}_bench_ref.json 2> gpurun_out/${T}_bench_ref.err; echo "ref rc=$?" >> gpurun_out/${T}_bench_ref.err
tail -n 3 gpurun_out/${T}_tests.log gpurun_out/${T}_smoke.log gpurun_out/${T}_bench.err gpurun_out/${T}_bench_ref.err
