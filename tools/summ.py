"""Print a one-line summary of bench.py JSON lines found in the given log files."""
import json
import sys

for path in sys.argv[1:]:
    for line in open(path):
        if line.startswith("{"):
            d = json.loads(line)
            k = d.get("kernels_us", {})
            print(f"{path}: n={d['n_gpus']} ms/step={d['ms_per_step']:.4f} e2e_us={d['latency_us']['e2e_median']:.1f} "
                  f"roll={k.get('rollout_cost', 0):.1f} upd={k.get('softmin_update', 0):.1f} fin={k.get('finalize_nominal', 0):.1f} "
                  f"fp32frac={d['roofline']['frac']:.3f} hbmfrac={d['roofline_update']['frac']:.3f} value={d['value']:.3e} e2e={d['e2e']['value']:.3e}")
