#!/bin/bash
# round-2 evidence, one call: GPU suite, smoke, default bench (+ reference arm), ncu launch list of the same command, full captures
# (CSV pages exported on the box) for cfg2 (default step and rollout_variant 7) / cfg3 / cfg4 (walls reached) / cfg5
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -n 3 gpurun_out/pytest_gpu.log
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -n 1
timeout 500 python bench.py > gpurun_out/bench_cfg2.json 2> gpurun_out/bench_cfg2.err; echo "bench rc=$?"
timeout 500 python bench.py --impl reference --steps 20 --warmup 3 > gpurun_out/bench_cfg2_reference.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 800 --csv --log-file gpurun_out/launches_cfg2.csv python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-extra > gpurun_out/ncu_l.log 2>&1
bash tools/r2_ncu.sh cfg2 "rollout_pipe_kernel|update_kernel|finalize_kernel|noise_kernel" 40 4 -- --steps 20 --warmup 3 --no-extra
MPPI_PIPE_FOLD=1 timeout 300 python bench.py --no-cpu-baseline --no-extra --steps 100 --warmup 10 > gpurun_out/bench_cfg2_variant7.json 2> gpurun_out/bench_cfg2_variant7.err; echo "bench v7 rc=$?"
bash tools/r2_ncu.sh cfg2v7 "rollout_pipe_kernel|finalize_kernel|noise_kernel" 30 3 MPPI_PIPE_FOLD=1 -- --steps 20 --warmup 3 --no-extra
bash tools/r2_ncu.sh cfg3 "rollout_kernel|update_kernel" 12 2 -- --config 2 --steps 3 --warmup 3
bash tools/r2_ncu.sh cfg4near "rollout_kernel|update_kernel" 12 2 -- --config 3 --x0 38,0,0,0,0,8,0 --steps 3 --warmup 3
bash tools/r2_ncu.sh cfg5 "rollout_kernel|update_kernel|finalize_kernel" 12 3 -- --config 4 --steps 3 --warmup 3
python tools/summ.py gpurun_out/bench_cfg2.json gpurun_out/bench_cfg2_variant7.json
