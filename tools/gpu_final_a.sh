#!/bin/bash
# GPU test suite + the default bench command (what the driver runs at round end).
tag=${1:-final}
mkdir -p gpurun_out
( time timeout 480 python -m pytest tests -m gpu -x -q ) > gpurun_out/${tag}_tests.log 2>&1
echo "tests rc=$?" >> gpurun_out/${tag}_tests.log
tail -n 5 gpurun_out/${tag}_tests.log
( time timeout 600 python bench.py > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err )
echo "bench rc=$?"
tail -n 4 gpurun_out/${tag}_bench.err
python - gpurun_out/${tag}_bench.json <<'PY'
import json, sys
d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print("ms/step %.1f value %.0f e2e %.0f" % (d["ms_per_step"], d["value"], d["e2e"]["value"]), {k: round(v, 1) for k, v in d["stage_ms"].items()})
print("roofline", d["roofline"]["kernel"], round(d["roofline"]["frac"], 4), "cpu", d.get("cpu_baseline", {}).get("value"), "e2e_sam", d.get("e2e_sam", {}).get("value"))
PY
