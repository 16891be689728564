#!/bin/bash
# round 2, call A: full GPU suite, default bench, scan-variant bench, ncu launch lists and full captures of both step forms
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -n 4 gpurun_out/pytest_gpu.log
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -n 2
timeout 400 python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; echo "bench rc=$?"
MPPI_ROLLOUT_VARIANT=5 timeout 400 python bench.py --no-cpu-baseline > gpurun_out/bench_scan.json 2> gpurun_out/bench_scan.err; echo "bench scan rc=$?"
for sc in 148 296 444 592; do
MPPI_SCAN_CTAS=$sc MPPI_ROLLOUT_VARIANT=5 timeout 400 python bench.py --no-cpu-baseline --steps 50 --warmup 5 > gpurun_out/bench_scan_$sc.json 2>/dev/null
done
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_cfg2.csv python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_l.log 2>&1
MPPI_ROLLOUT_VARIANT=5 timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_cfg2_scan.csv python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_ls.log 2>&1
MPPI_ROLLOUT_VARIANT=5 timeout 400 ncu --set full --clock-control none --import-source on -k regex:"scan_rollout_kernel" -s 30 -c 1 -o gpurun_out/prof_scan python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_fs.log 2>&1
timeout 400 ncu --set full --clock-control none --import-source on -k regex:"rollout_kernel|update_kernel|finalize_kernel" -s 40 -c 3 -o gpurun_out/prof_seq python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_f2.log 2>&1
python tools/summ.py gpurun_out/bench_default.json gpurun_out/bench_scan.json gpurun_out/bench_scan_*.json
ls -la gpurun_out/
