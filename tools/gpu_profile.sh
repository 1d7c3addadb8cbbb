#!/bin/bash
# ncu evidence for the bench command on one B200 (run under gpurun, after the same command has exited 0 without
# ncu): every launch of one run with its device time and DRAM bytes (-> tools/launch_table.py, tools/make_traffic.py),
# then a full capture with source of the kernels whose demangled name matches a regex (-> ncu --page raw --csv,
# tools/ncu_hotspots.py).  Steady-state launches of the search / screening kernels start at the 7th match
# (the first batch of a context sizes its lists with a few voided attempts).
# usage: tools/gpu_profile.sh <tag> <kernel regex> <skip> <count> [reads per step]
tag=$1; pat=$2; skip=${3:-6}; cnt=${4:-3}; reads=${5:-1000000}
mkdir -p gpurun_out
CMD="python bench.py --reads $reads --steps 1 --warmup 3 --no-cpu-baseline --no-e2e-sam --host-threads 1"
timeout 300 $CMD > gpurun_out/${tag}_plain.log 2> gpurun_out/${tag}_plain.err || { echo "plain run failed"; exit 1; }
timeout 300 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 4000 --csv \
    --log-file gpurun_out/${tag}_launches.csv $CMD > gpurun_out/${tag}_ncu1.log 2>&1
echo "launch list rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k "regex:$pat" -s $skip -c $cnt \
    -o gpurun_out/${tag}_full $CMD > gpurun_out/${tag}_ncu2.log 2>&1
echo "full capture rc=$?"
tail -n 3 gpurun_out/${tag}_ncu2.log | cut -c1-300
