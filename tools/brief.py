"""Print the few numbers of a bench.py JSON line that matter while tuning (reads stdin)."""
import json, sys
tag = sys.argv[1] if len(sys.argv) > 1 else ""
d = json.loads(sys.stdin.read().strip().splitlines()[-1])
r = d.get("roofline", {})
print(tag, f"value {d['value']/1e6:.1f} M/s  step {d['ms_per_step']:.4f} ms  K1 {r.get('kernel_ms', 0):.4f} ms  "
           f"frac {r.get('frac', 0):.3f}  e2e {d['e2e']['value']/1e6:.1f} M/s  launches {d['gpu_launches']}")
