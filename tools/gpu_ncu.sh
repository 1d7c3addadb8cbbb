#!/bin/bash
# One gpurun call: ncu --set full capture (with source) of one steady-state launch of the two search kernels and
# the screening kernel of the bench command.  usage: tools/gpu_ncu.sh <tag> [skip] [count]
tag=${1:-ncu}; skip=${2:-6}; cnt=${3:-3}
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-e2e-sam --host-threads 1"
timeout 300 $CMD > gpurun_out/${tag}_plain.json 2> gpurun_out/${tag}_plain.err || { echo "plain run failed"; exit 1; }
timeout 600 ncu --set full --clock-control none --import-source on --kernel-name-base demangled \
    -k "regex:FrontierKernel|ScreenKernel" -s $skip -c $cnt -o gpurun_out/${tag}_full $CMD > gpurun_out/${tag}_ncu.log 2>&1
echo "ncu rc=$?"
ls -la gpurun_out/${tag}_full.ncu-rep
