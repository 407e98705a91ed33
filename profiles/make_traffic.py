#!/usr/bin/env python
"""Regenerate profiles/traffic.json (the `traffic` field of bench.py's roofline entries) from a summary written by
profiles/ncu_summary.py: dram__bytes_read.sum + dram__bytes_write.sum per launch, summed over the kernels of a stage,
for the first complete view (preprocess ... preprocess backward) in the capture.
Usage: python profiles/make_traffic.py profiles/rNN_ncu_all_kernels.txt"""
import json
import os
import re
import sys

src = sys.argv[1]
UNIT = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
launches = []  # (kernel name, dram bytes)
name, rd = None, 0.0
for line in open(src):
    if line.startswith("=="):
        name, rd = line[2:].strip(), 0.0
    m = re.match(r"\s+dram__bytes_(read|write)\.sum\s+([\d.]+)\s+(\w+)", line)
    if m:
        b = float(m.group(2)) * UNIT[m.group(3)]
        if m.group(1) == "read":
            rd = b
        else:
            launches.append((name, rd + b))

stage_of = {"preprocess_kernel": "preprocess", "scan_kernel": "scan", "duplicate_kernel": "duplicate", "ranges_kernel": "ranges",
            "render_forward_kernel": "blend_fwd", "image_loss_kernel": "loss", "render_backward_kernel": "blend_bwd",
            "preprocess_backward": "preprocess_bwd"}
out, started, after_duplicate = {}, False, False
for name, b in launches:
    if "preprocess_kernel" in name:
        if started:
            break
        started = True
    if not started:
        continue
    if "sort_" in name:
        st = "sort" if after_duplicate else "depth_sort"
    else:
        st = next((v for k, v in stage_of.items() if k in name), None)
        if st is None:
            continue
        after_duplicate |= st == "duplicate"
    out[st] = out.get(st, 0) + int(round(b))
    if st == "preprocess_bwd":
        break
doc = {"_source": "%s: dram__bytes_read.sum + dram__bytes_write.sum per launch, ncu --set full, one C4 view (3M Gaussians, "
                  "1600x1064); multi-kernel stages are summed over their kernels (profiles/make_traffic.py)" % src}
try:  # the commit the capture was taken at (the working tree's HEAD when this script runs right after the capture)
    import subprocess
    doc["_commit"] = subprocess.run(["git", "rev-parse", "--short", "HEAD"], capture_output=True, text=True,
                                    cwd=os.path.dirname(os.path.abspath(__file__))).stdout.strip() or None
except OSError:
    doc["_commit"] = None
doc["_file"] = os.path.basename(src)
doc.update(out)
json.dump(doc, open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "traffic.json"), "w"), indent=1)
print(json.dumps(doc, indent=1))
