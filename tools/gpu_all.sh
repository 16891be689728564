#!/bin/bash
# parity tests + a short bench of every BASELINE.json config on one GPU
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -n 3 gpurun_out/pytest_gpu.log
for c in 1 2 3 4; do
  timeout 280 python bench.py --config $c --steps 30 --warmup 5 --no-cpu-baseline > gpurun_out/bench_c$c.log 2>&1 || { echo "cfg $c FAILED"; tail -n 5 gpurun_out/bench_c$c.log; }
done
python tools/summ.py gpurun_out/bench_c1.log gpurun_out/bench_c2.log gpurun_out/bench_c3.log gpurun_out/bench_c4.log
