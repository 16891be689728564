#!/bin/bash
# Round-end evidence. part 1: headline config (cfg2): plain bench, ncu launch list of the same command,
# full capture of the step kernels.  parts 2 and 3: the large configs (cfg3, cfg4): rollout + update kernels.
set -x
if [ "$1" = "1" ]; then
python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/plain_cfg2.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_cfg2.csv python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_l.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:"rollout_kernel|update_kernel|finalize_kernel|noise_kernel" -s 40 -c 4 -o gpurun_out/prof_cfg2 python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_f2.log 2>&1
else
for c in $(( $1 )); do   # 2 = cfg3, 3 = cfg4: one config per call, the reports are ~30 MB each and gpurun_out/ is capped at 64 MiB
python bench.py --config $c --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/plain_c$c.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:"rollout_kernel|update_kernel" -s 12 -c 2 -o gpurun_out/prof_c$c python bench.py --config $c --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_f$c.log 2>&1
done
fi
ls -la gpurun_out/
