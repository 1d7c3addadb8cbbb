#!/bin/bash
# One gpurun call: short bench runs of the prefetch / 32-bit verification variants, then one ncu --set full capture
# (with source) of a steady-state launch of the two search kernels and the screening kernel.
# usage: tools/gpu_ab2.sh <tag>
tag=${1:-ab2}
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
    print(sys.argv[1], "ms/step %.1f" % d["ms_per_step"], "e2e %.0f" % d["e2e"]["value"], {k: round(v, 1) for k, v in d["stage_ms"].items()})
except Exception as e:
    print(sys.argv[1], "no line:", e)
PY
}
run p0_v64 COLUMBA_B200_SEARCH_PREFETCH=0 COLUMBA_B200_VERIFY32=0
run p1_v64 COLUMBA_B200_SEARCH_PREFETCH=1 COLUMBA_B200_VERIFY32=0
run p1_v32 COLUMBA_B200_SEARCH_PREFETCH=1 COLUMBA_B200_VERIFY32=1
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-e2e-sam --host-threads 1"
timeout 600 ncu --set full --clock-control none --import-source on --kernel-name-base demangled \
    -k "regex:FrontierKernel|ScreenKernel" -s 6 -c 3 -o gpurun_out/${tag}_full $CMD > gpurun_out/${tag}_ncu.log 2>&1
echo "ncu rc=$?"
tail -n 3 gpurun_out/${tag}_ncu.log | cut -c1-300
ls -la gpurun_out/${tag}_full.ncu-rep
