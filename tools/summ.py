"""Print a one-line summary of bench.py JSON lines found in the given log files."""
import json
import sys

for path in sys.argv[1:]:
    try:
        lines = open(path).read().splitlines()
    except OSError:
        continue
    for line in lines:
        if line.startswith("{"):
            d = json.loads(line)
            if 'roofline' not in d:   # the reference arm's line
                print(f"{path}: impl={d.get('impl')} ms/step={d.get('ms_per_step', 0):.3f} value={d.get('value', 0):.3e}")
                continue
            k = d.get("kernels_us", {})
            g = d.get("kernels_us_in_graph") or {}
            lat = d.get("latency_us", {})
            ru = d.get("roofline_update") or {}
            print(f"{path}: n={d['n_gpus']} ms/step={d['ms_per_step']:.4f} dev_med={lat.get('device_median', 0):.1f} p99={lat.get('device_p99', 0):.1f} "
                  f"e2e_res={lat.get('e2e_result_median', 0):.1f} e2e_idle={lat.get('e2e_idle_median', 0):.1f} "
                  f"roll={k.get('rollout_cost', 0):.1f} upd={k.get('softmin_update', 0):.1f} fin={k.get('finalize_nominal', 0):.1f} "
                  f"| graph roll={g.get('rollout_cost', 0):.1f} upd={g.get('softmin_update', 0):.1f} fin={g.get('finalize_nominal', 0):.1f} "
                  f"span={g.get('rollout_start_to_finalize_end', 0):.1f} "
                  f"| fp32frac={d['roofline']['frac']:.3f} ingraph={d['roofline'].get('frac_in_graph') or 0:.3f} hbmfrac={ru.get('frac') or 0:.3f} "
                  f"value={d['value']:.3e} e2e={d['e2e']['value']:.3e} launches={d.get('gpu_launches')}"
                  + (f" | c_abi res={d['e2e_c_abi']['result_median_us']:.1f} idle={d['e2e_c_abi']['idle_median_us']:.1f} value={d['e2e_c_abi']['value']:.3e}" if d.get('e2e_c_abi') else ""))
