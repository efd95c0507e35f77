"""Print the headline numbers and the per-stage table of bench.py JSON lines (one file per argument)."""
import json
import sys

for path in sys.argv[1:]:
    for line in open(path):
        if not line.startswith("{"):
            continue
        d = json.loads(line)
        e2e = d.get("e2e") or {}
        print(f"{path}: {d.get('value', 0):.1f} {d.get('unit')}  {d.get('ms_per_step', 0):.4f} ms/step  e2e {e2e.get('ms_per_step', 0):.4f} ms"
              f"  raster {d.get('raster_fwd_bwd_ms')}")
        st = d.get("stages") or {}
        print("   " + "  ".join(f"{k}={v['ms']:.3f}" for k, v in st.items()))
        for key in ("sustained", "best_case", "stress", "north_star_batch"):
            v = d.get(key)
            if v:
                print(f"   {key}: {v.get('ms_per_step', 0):.4f} ms/step  " + json.dumps({k: v[k] for k in v if k not in ('clocks',)})[:600])
        try:
            print("   roofline:", json.dumps(d.get("roofline")), " idle", d.get("adam_idle_frac"), " clocks", d.get("clocks"))
        except BrokenPipeError:
            pass
