#!/bin/bash
# ncu full captures exported to CSV on the box (the .ncu-rep files carry the 17 MB cubin and exceed the 64 MiB return cap)
# usage: r2_ncu.sh <tag> <kernel-regex> <skip> <count> [env assignments...] -- bench args
tag=$1; rx=$2; skip=$3; cnt=$4; shift 4
envs=()
while [ "$1" != "--" ] && [ $# -gt 0 ]; do envs+=("$1"); shift; done
shift
mkdir -p gpurun_out
env "${envs[@]}" timeout 600 ncu --set full --clock-control none --import-source on -k regex:"$rx" -s $skip -c $cnt -o /tmp/prof_$tag python bench.py "$@" --no-cpu-baseline > gpurun_out/ncu_$tag.log 2>&1
ncu -i /tmp/prof_$tag.ncu-rep --page raw --csv > gpurun_out/ncu_${tag}_raw.csv 2>/dev/null
ncu -i /tmp/prof_$tag.ncu-rep --page source --csv > gpurun_out/ncu_${tag}_source.csv 2>/dev/null
ls -la gpurun_out/ncu_${tag}_*
