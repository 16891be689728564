#!/bin/bash
# quick cycle: selected tests, then the default bench line
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q ${1:+-k "$1"} > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -n 25 gpurun_out/pytest_gpu.log
timeout 300 python bench.py --no-cpu-baseline > gpurun_out/bench_q.json 2> gpurun_out/bench_q.err; echo "bench rc=$?"; tail -c 600 gpurun_out/bench_q.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_q.json').read().strip().splitlines()[-1])
print({k:d[k] for k in ('ms_per_step','gpu_launches','kernels_us','kernels_us_in_graph','latency_us')})
print('e2e',d['e2e']['value'],'roofline',d['roofline']['frac'],d['roofline'].get('frac_in_graph'))
PY
