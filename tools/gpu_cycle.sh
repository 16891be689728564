#!/bin/bash
# one GPU cycle: parity tests, default bench, full ncu capture of the three step kernels
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
timeout 300 python bench.py --no-cpu-baseline > gpurun_out/bench.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:"rollout_kernel|update_kernel|finalize_kernel" -s 24 -c 3 -o gpurun_out/prof_cycle python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/ncu2.log 2>&1
tail -n 4 gpurun_out/pytest_gpu.log; cat gpurun_out/bench.log
