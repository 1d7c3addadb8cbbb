#!/bin/bash
# ncu evidence for the bench command on one B200 (run under gpurun): every launch of the run with its device
# time and DRAM bytes, then a full capture of the kernels whose demangled name matches a regex.
# usage: tools/gpu_profile.sh <tag> <kernel regex> <skip> <count> [reads per step]
tag=$1; pat=$2; skip=${3:-0}; cnt=${4:-4}; reads=${5:-1000000}
CMD="python bench.py --reads $reads --steps 1 --warmup 3 --no-cpu-baseline --host-threads 1"
$CMD > gpurun_out/${tag}_plain.log 2> gpurun_out/${tag}_plain.err &&
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 4000 --csv \
    --log-file gpurun_out/${tag}_launches.csv $CMD > gpurun_out/${tag}_ncu1.log 2>&1
$CMD > /dev/null 2>&1 &&
ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k "regex:$pat" -s $skip -c $cnt \
    -o gpurun_out/${tag}_full $CMD > gpurun_out/${tag}_ncu2.log 2>&1
tail -n 3 gpurun_out/${tag}_ncu2.log | cut -c1-300
