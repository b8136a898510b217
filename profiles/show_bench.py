"""Print a one-line summary per bench JSON file (gpurun_out/bench_*.json)."""
import json, sys
for f in sys.argv[1:]:
    try:
        d = json.loads([l for l in open(f).read().splitlines() if l.startswith('{')][-1])
    except Exception as e:
        print(f, "unreadable", e); continue
    r = d.get("roofline") or {}
    e = d.get("e2e") or {}
    print(f"{f.split('/')[-1]:22s} n={d.get('n_gpus')} ms/step {d['ms_per_step']:.4f} planes/s {d['value']:.3e} "
          f"GB/s {r.get('achieved', 0):7.1f} frac {r.get('frac', 0):.3f} e2e_ms {e.get('ms_per_step', 0):.4f} "
          f"ceil {({k: int(v) for k, v in (r.get('write_only_ceilings') or {}).items()})} clk {d.get('clocks', {}).get('sm_mhz') if d.get('clocks') else None}")
