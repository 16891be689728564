#!/bin/bash
mkdir -p gpurun_out
for i in "$@"; do
  MPPI_B200_LIB=build_exp/lib_$i.so timeout 300 python bench.py --no-cpu-baseline --no-extra --steps 100 --warmup 10 > gpurun_out/bench_exp_$i.json 2> gpurun_out/bench_exp_$i.err || tail -3 gpurun_out/bench_exp_$i.err
done
python tools/summ.py gpurun_out/bench_exp_*.json
