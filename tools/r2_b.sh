#!/bin/bash
mkdir -p gpurun_out
timeout 400 python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; echo "bench rc=$?"
MPPI_ROLLOUT_VARIANT=5 timeout 400 python bench.py --no-cpu-baseline > gpurun_out/bench_scan.json 2> gpurun_out/bench_scan.err; echo "bench scan rc=$?"
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_cfg2.csv python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_l.log 2>&1
MPPI_ROLLOUT_VARIANT=5 timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_cfg2_scan.csv python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_ls.log 2>&1
bash tools/r2_ncu.sh scan "scan_rollout_kernel" 30 1 MPPI_ROLLOUT_VARIANT=5 -- --steps 20 --warmup 3
bash tools/r2_ncu.sh seq "rollout_kernel|update_kernel|finalize_kernel" 40 3 -- --steps 20 --warmup 3
python tools/summ.py gpurun_out/bench_default.json gpurun_out/bench_scan.json
