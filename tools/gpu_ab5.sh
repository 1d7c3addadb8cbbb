#!/bin/bash
tag=${1:-ab5}
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
run rp0 COLUMBA_B200_RANK_PREFETCH=0
run rp1 COLUMBA_B200_RANK_PREFETCH=1
run rp1_hw32k COLUMBA_B200_RANK_PREFETCH=1 COLUMBA_B200_HEAVY_WIDTH=32768
