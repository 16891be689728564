for i in 1 2 3; do
timeout 300 python bench.py --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
g=d['kernels_us_in_graph']
print(round(d['ms_per_step']*1e3,2), d['latency_us']['device_median'], 'e2e', round(d['latency_us']['e2e_median'],2), 'roll', g['rollout_cost'], 'fin_end', g['rollout_start_to_finalize_end'], 'last', g['rollout_start_to_last_kernel_end'])"
done
