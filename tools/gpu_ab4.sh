#!/bin/bash
# One gpurun call: short bench runs of search-stage variants (heavy-width shortcut, occupancy, scratch size).
tag=${1:-ab4}
mkdir -p gpurun_out
B="python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-e2e-sam"
run() { # name, env...
  name=$1; shift
  env "$@" timeout 300 $B > gpurun_out/${tag}_bench_$name.json 2> gpurun_out/${tag}_bench_$name.err
  echo "$name rc=$?"
  python - "$name" gpurun_out/${tag}_bench_$name.json <<'PY'
import json, sys
try:
    d = json.loads(open(sys.argv[2]).read().strip().splitlines()[-1])
    print(sys.argv[1], "ms/step %.1f" % d["ms_per_step"], "e2e %.0f" % d["e2e"]["value"], {k: round(v, 1) for k, v in d["stage_ms"].items()}, "heavy", d["stats_per_step"].get("HEAVY_TASKS"))
except Exception as e:
    print(sys.argv[1], "no line:", e)
PY
}
run base A=1
run hw1000 COLUMBA_B200_HEAVY_WIDTH=1000
run hw4000 COLUMBA_B200_HEAVY_WIDTH=4000
run hw16000 COLUMBA_B200_HEAVY_WIDTH=16000
run occ4 COLUMBA_B200_SEARCH_OCC=4
run slab512k COLUMBA_B200_SEARCH_SLAB=524288
