#!/bin/bash
# One gpurun call: GPU test suite, then short bench runs of the heavy-launch occupancy variants.
tag=${1:-ab3}
mkdir -p gpurun_out
( time timeout 480 python -m pytest tests -m gpu -x -q ) > gpurun_out/${tag}_tests.log 2>&1
echo "tests rc=$?" >> gpurun_out/${tag}_tests.log
tail -n 5 gpurun_out/${tag}_tests.log
B="python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-e2e-sam"
run() { # name, env...
  name=$1; shift
  env "$@" timeout 300 $B > gpurun_out/${tag}_bench_$name.json 2> gpurun_out/${tag}_bench_$name.err
  echo "$name rc=$?"
  python - "$name" gpurun_out/${tag}_bench_$name.json <<'PY'
import json, sys
try:
    d = json.loads(open(sys.argv[2]).read().strip().splitlines()[-1])
    print(sys.argv[1], "ms/step %.1f" % d["ms_per_step"], "e2e %.0f" % d["e2e"]["value"], {k: round(v, 1) for k, v in d["stage_ms"].items()})
except Exception as e:
    print(sys.argv[1], "no line:", e)
PY
}
run h512 COLUMBA_B200_HEAVY_PER_SM=512
run h1024 COLUMBA_B200_HEAVY_PER_SM=1024
run h1024_g256 COLUMBA_B200_HEAVY_PER_SM=1024 COLUMBA_B200_HEAVY_GROUP=256
