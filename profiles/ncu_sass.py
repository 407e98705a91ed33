#!/usr/bin/env python
"""Per-SASS-instruction executed counts and stall samples of one kernel in an .ncu-rep.
Usage: python profiles/ncu_sass.py rep.ncu-rep <kernel regex> [min_exec_millions]"""
import csv, subprocess, sys, io
rep, kre = sys.argv[1], sys.argv[2]
thr = float(sys.argv[3]) if len(sys.argv) > 3 else 0.0
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", "regex:" + kre],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
# first kernel only
hdr = rows[1]
iS, iE, iN, iT = hdr.index("Source"), hdr.index("Instructions Executed"), hdr.index("# Samples"), hdr.index("Avg. Threads Executed")
stall_cols = [(i, h) for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
tot = 0; tots = 0
out = []
for r in rows[2:]:
    if len(r) < len(hdr) or r[0] == "Address":
        if r and r[0] == "Kernel Name": break
        continue
    e = int(r[iE]); n = int(r[iN]); tot += e; tots += n
    st = sorted(((int(r[i]), h[6:]) for i, h in stall_cols if r[i].isdigit() and int(r[i]) > 0), reverse=True)[:2]
    out.append((e, n, r[iT], r[iS].strip(), st))
print(f"total executed {tot/1e6:.1f} M warp-instr, samples {tots}")
for k, (e, n, t, s, st) in enumerate(out):
    if e / 1e6 >= thr:
        print(f"{k:4d} {e/1e6:8.2f}M {n:6d} thr={t:>5s} {s[:70]:70s} {' '.join(f'{h}:{v}' for v,h in st)}")
