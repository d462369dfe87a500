"""Print the headline fields of a bench.py JSON line (development aid): python scripts/show_bench.py FILE"""
import json
import sys

for line in open(sys.argv[1]):
    if not line.startswith("{"):
        print("EXTRA LINE:", line[:100].rstrip())
        continue
    d = json.loads(line)
    sm = d.get("sub_metrics", {})
    print("n_gpus", d.get("n_gpus"), "value", d.get("value"), "e2e", d.get("e2e", {}).get("value"), "verified", sm.get("verified", sm.get("verified_known_dlog")))
    print("stage_ms", {k: round(v, 2) for k, v in sm.get("stage_ms", {}).items()})
    print("kernel_ms", {k: round(v, 2) for k, v in sm.get("kernel_ms_per_step", {}).items()})
    print("roofline frac", d.get("roofline", {}).get("frac"), "executed", d.get("roofline", {}).get("executed_frac"), "clocks", d.get("clocks"))
