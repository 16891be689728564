#!/bin/bash
# full GPU suite + default bench line + the same with the pipeline's in-block update (rollout_variant 7 through MPPI_PIPE_FOLD=1)
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_gpu.log
timeout 300 python bench.py --no-cpu-baseline --steps 100 --warmup 10 > gpurun_out/bench_cfg2_v6.json 2> gpurun_out/bench_cfg2_v6.err || tail -3 gpurun_out/bench_cfg2_v6.err
MPPI_PIPE_FOLD=1 timeout 300 python bench.py --no-cpu-baseline --no-extra --steps 100 --warmup 10 > gpurun_out/bench_cfg2_v7.json 2> gpurun_out/bench_cfg2_v7.err || tail -3 gpurun_out/bench_cfg2_v7.err
python tools/summ.py gpurun_out/bench_cfg2_v6.json gpurun_out/bench_cfg2_v7.json
