#!/bin/bash
# A/B runs of the bench command on one B200 (run under gpurun): every argument is one variant,
#   name[:VAR=value[,VAR=value...]]
# run back to back on the same box with the measurement switches of DESIGN.md section 3 in its environment.
# Prints one line per variant (ms per step, end-to-end reads/s, stage times) and leaves the full bench lines in
# gpurun_out/<tag>_bench_<name>.json.  With TESTS=1 the GPU test suite runs first.
# usage: [TESTS=1] [TAG=r2s] tools/gpu_ab.sh g32:COLUMBA_B200_SEARCH_GROUP=32 g16:COLUMBA_B200_SEARCH_GROUP=16 ...
tag=${TAG:-ab}
mkdir -p gpurun_out
if [ -n "$TESTS" ]; then
  ( time timeout 480 python -m pytest tests -m gpu -x -q ) > gpurun_out/${tag}_tests.log 2>&1
  echo "tests rc=$?" >> gpurun_out/${tag}_tests.log
  tail -n 5 gpurun_out/${tag}_tests.log
fi
B="python bench.py --steps ${STEPS:-5} --warmup 3 --no-cpu-baseline --no-e2e-sam"
for spec in "$@"; do
  name=${spec%%:*}
  envs=""
  [ "$spec" != "$name" ] && envs=$(echo "${spec#*:}" | tr ',' ' ')
  env $envs timeout 300 $B > gpurun_out/${tag}_bench_$name.json 2> gpurun_out/${tag}_bench_$name.err
  echo "$name rc=$?"
  python - "$name" gpurun_out/${tag}_bench_$name.json <<'PY'
import json, sys
try:
    d = json.loads(open(sys.argv[2]).read().strip().splitlines()[-1])
    print(sys.argv[1], "ms/step %.1f" % d["ms_per_step"], "e2e %.0f" % d["e2e"]["value"],
          {k: round(v, 1) for k, v in d["stage_ms"].items()}, "heavy", d["stats_per_step"].get("HEAVY_TASKS"))
except Exception as e:
    print(sys.argv[1], "no line:", e)
PY
done
