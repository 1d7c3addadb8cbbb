#!/bin/bash
# One gpurun call: the GPU test suite, the 3.1 Gbp parity test again with half / quarter-warp search groups, then
# short bench runs of the search-group and screening variants (device-resident value + stage times).
# usage: tools/gpu_ab.sh <tag>
tag=${1:-ab}
mkdir -p gpurun_out
( time timeout 420 python -m pytest tests -m gpu -x -q ) > gpurun_out/${tag}_tests.log 2>&1
echo "tests rc=$?" >> gpurun_out/${tag}_tests.log
tail -n 4 gpurun_out/${tag}_tests.log
for g in 8 16; do
  ( time COLUMBA_B200_SEARCH_GROUP=$g timeout 300 python -m pytest tests/test_gpu_human3g.py tests/test_gpu_parity.py -m gpu -x -q ) > gpurun_out/${tag}_tests_g$g.log 2>&1
  echo "tests g$g rc=$?" >> gpurun_out/${tag}_tests_g$g.log
  tail -n 3 gpurun_out/${tag}_tests_g$g.log
done
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
run g32_s64 COLUMBA_B200_SEARCH_GROUP=32 COLUMBA_B200_SCREEN32=0
run g32_s32 COLUMBA_B200_SEARCH_GROUP=32 COLUMBA_B200_SCREEN32=1
run g16_s32 COLUMBA_B200_SEARCH_GROUP=16 COLUMBA_B200_SCREEN32=1
run g8_s32 COLUMBA_B200_SEARCH_GROUP=8 COLUMBA_B200_SCREEN32=1
