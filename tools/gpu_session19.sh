#!/bin/bash
# K x share retune after the round's kernel changes (spare SMs go to the first pipelines)
mkdir -p gpurun_out
for ks in ${SWEEP:-16:9 18:8 14:10 12:12 21:7 13:11 20:7 24:6}; do
  K=${ks%%:*}; S=${ks##*:}
  python bench.py --no-cpu-baseline --no-direction-sharded --concurrency $K --sm-share $S > gpurun_out/r2s52_${K}x${S}.json 2>/dev/null
  python - <<PY
import json
for l in open("gpurun_out/r2s52_${K}x${S}.json"):
    if l.startswith("{"):
        d=json.loads(l); print("K=$K share=$S: %.0f frames/s resident, %.0f e2e, mismatches %d, +1 SM on %d pipelines" % (d["value"], d["e2e"]["value"], d["parity_in_timed_region"]["mismatches"], d["config"]["pipelines_with_one_more_sm"]))
PY
done
