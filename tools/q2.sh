for v in "MPPI_ROLLOUT_VARIANT=1" "MPPI_ROLLOUT_VARIANT=2 MPPI_PREGEN_PACKED=1" "MPPI_ROLLOUT_VARIANT=2"; do
  echo "== $v"
  env $v timeout 300 python bench.py --no-cpu-baseline --steps 100 --warmup 10 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print(d['ms_per_step'], d['gpu_launches'], d['kernels_us_in_graph'], d['latency_us'])"
done
