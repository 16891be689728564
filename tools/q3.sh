for c in 2 3; do
for v in "MPPI_PREGEN=0" "MPPI_PREGEN_PACKED=1"; do
  echo "== cfg index $c $v"
  env $v timeout 300 python bench.py --config $c --no-cpu-baseline --steps 30 --warmup 5 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print(d['ms_per_step'], d['gpu_launches'], {k:v for k,v in d['kernels_us_in_graph'].items() if k!='source'}, d['latency_us']['e2e_median'])"
done
done
