"""Condense bench.py JSON lines (stdin) to one short line each."""
import json, sys
for line in sys.stdin:
    line = line.strip()
    if not line.startswith("{"):
        continue
    d = json.loads(line)
    c, r, e = d.get("config", {}), d.get("roofline", {}), d.get("e2e", {})
    print(f"skip={c.get('skip_block')} exact={c.get('exact_filter')} tfsmem={c.get('tf_in_smem')} n={d.get('n_gpus')} "
          f"ms={d.get('ms_per_step'):.4f} fps={c.get('fps'):.0f} kernel_ms={r.get('kernel_ms', 0):.4f} fetched={c.get('samples_per_frame')} "
          f"texfrac={r.get('frac', 0):.3f} e2e_ms={e.get('ms_per_step', 0):.3f} value={d.get('value'):.4g}")
