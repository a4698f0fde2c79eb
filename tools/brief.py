"""Print the few numbers of a bench.py JSON line that matter while tuning: brief.py FILE [TAG] (or TAG < FILE)."""
import json, os, sys
args = sys.argv[1:]
if args and os.path.isfile(args[0]):
    text, tag = open(args[0]).read(), (args[1] if len(args) > 1 else "")
else:
    text, tag = sys.stdin.read(), (args[0] if args else "")
d = json.loads(text.strip().splitlines()[-1])
r = d.get("roofline", {})
print(tag, f"value {d['value']/1e6:.1f} M/s  step {d['ms_per_step']:.4f} ms  K1 {r.get('kernel_ms', 0):.4f} ms  "
           f"frac {r.get('frac', 0):.3f}  e2e {d['e2e']['value']/1e6:.1f} M/s  launches {d['gpu_launches']}")
