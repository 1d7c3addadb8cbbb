#!/bin/bash
# Short bench of the final build, ncu launch list of one bench run (device time + DRAM bytes per launch), then an
# ncu --set full capture of one steady-state launch of the two search kernels.
tag=${1:-finalb}
mkdir -p gpurun_out
timeout 200 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-e2e-sam > gpurun_out/${tag}_bench_short.json 2> gpurun_out/${tag}_bench_short.err
echo "bench rc=$?"
python - gpurun_out/${tag}_bench_short.json <<'PY'
import json, sys
d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print("ms/step %.1f value %.0f e2e %.0f" % (d["ms_per_step"], d["value"], d["e2e"]["value"]), {k: round(v, 1) for k, v in d["stage_ms"].items()})
PY
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-e2e-sam --host-threads 1"
( time timeout 200 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 4000 --csv \
    --log-file gpurun_out/${tag}_launches.csv $CMD > gpurun_out/${tag}_ncu1.log 2>&1 )
echo "launch list rc=$?"
( time timeout 150 ncu --set full --clock-control none --import-source on --kernel-name-base demangled \
    -k "regex:FrontierKernel" -s 4 -c 2 -o gpurun_out/${tag}_full $CMD > gpurun_out/${tag}_ncu2.log 2>&1 )
echo "ncu full rc=$?"
ls -la gpurun_out/${tag}_full.ncu-rep gpurun_out/${tag}_launches.csv
