#!/bin/bash
# SURVEY 8(d) measurement set from the current code: cfg3, cfg4 (M_p = 3 / 16, default start and a start from which the walls are
# reached), cfg5; plain bench lines into gpurun_out/, then ncu (rollout + update kernels) exported to CSV on the box
mkdir -p gpurun_out
run() { tag=$1; shift; timeout 600 python bench.py "$@" --steps 30 --warmup 5 > gpurun_out/bench_$tag.json 2> gpurun_out/bench_$tag.err || { echo "$tag FAILED"; tail -3 gpurun_out/bench_$tag.err; }; }
run cfg3 --config 2
run cfg4 --config 3
run cfg4_mp16 --config 3 --walls 16
run cfg4_near --config 3 --x0 38,0,0,0,0,8,0
run cfg4_mp16_near --config 3 --walls 16 --x0 38,0,0,0,0,8,0
run cfg5 --config 4
python tools/summ.py gpurun_out/bench_cfg*.json
python - <<'PY'
import json, glob
for f in sorted(glob.glob('gpurun_out/bench_cfg*.json')):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
    except Exception as e:
        print(f, 'unreadable', e); continue
    r = d['roofline']; u = d.get('roofline_update') or {}
    print(f, 'sat_exact', r.get('sat_exact_fraction'), 'frac', r.get('frac'), 'frac_exec', r.get('frac_executed'), 'upd', u.get('frac'), u.get('frac_of_nominal_8TBs'),
          'cpu', (d.get('cpu_baseline') or {}).get('value'), 'cpu1', ((d.get('cpu_baseline') or {}).get('single_thread') or {}).get('value'))
PY
