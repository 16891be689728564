#!/bin/bash
mkdir -p gpurun_out
export MPPI_B200_LIB=build_exp/lib_0.so
MPPI_NOISE_CHAIN=1 timeout 600 python -m pytest tests/test_gpu_pipe.py tests/test_gpu_scan.py -m gpu -x -q 2>&1 | tail -2
for c in 0 1 0 1; do
  MPPI_NOISE_CHAIN=$c timeout 300 python bench.py --no-cpu-baseline --no-extra --steps 200 --warmup 20 > gpurun_out/bench_chain_$c.json 2> gpurun_out/bench_chain_$c.err || tail -3 gpurun_out/bench_chain_$c.err
  echo "chain=$c"; python tools/summ.py gpurun_out/bench_chain_$c.json
done
MPPI_NOISE_CHAIN=1 timeout 100 python tools/dbg/overhead2.py 2>&1 | tail -4
