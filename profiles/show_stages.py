#!/usr/bin/env python
"""Print the per-stage table of a bench.py JSON line.  Usage: python profiles/show_stages.py <log>"""
import json, sys
d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print(f"value {d['value']:.1f} {d['unit']}  ms/view/gpu {d['ms_per_view_per_gpu']:.3f}  e2e {d['e2e']['value'] if d.get('e2e') else None}")
tot = 0
for s in d["stages"]:
    tot += s["avg_ms"]
    print(f"{s['stage']:16s} {s['avg_ms']:.4f} ms  share {s['share']:.3f}  {s['achieved']:9.1f} {s['unit']:8s} frac {s['frac']:.3f}")
print(f"sum of stages {tot:.3f} ms;", d.get("work_per_view"))
