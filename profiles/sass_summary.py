#!/usr/bin/env python
"""Per-kernel SASS mnemonic summary of lib/libpdgs.so (cuobjdump -sass, no GPU needed): instruction count and the
mnemonics that show which hardware paths a kernel uses.  Usage: python profiles/sass_summary.py > profiles/rNN_sass_summary.txt"""
import collections
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "priordepth-gaussiansplatting_b200", "lib", "libpdgs.so")
SHOW = ("BAR", "FADD", "FFMA", "FMUL", "LDG", "LDS", "STG", "STS", "MUFU.EX2", "MUFU.RCP", "MUFU.RSQ", "REDG", "ATOMG", "ATOMS",
        "UBLKCP", "SYNCS", "MATCH", "VOTE", "FLO", "BMSK", "SHFL", "LDGSTS", "UTMALDG", "UTCMMA", "LDTM", "HMMA")
out = subprocess.run(["/usr/local/cuda/bin/cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
head = subprocess.run(["git", "rev-parse", "--short", "HEAD"], capture_output=True, text=True, cwd=ROOT).stdout.strip()
print("# cuobjdump -sass of lib/libpdgs.so built from the working tree at %s: per kernel, instruction count and the mnemonics" % head)
print("# that show which hardware paths the kernel uses (REDG = fire-and-forget global reductions, UBLKCP/SYNCS = cp.async.bulk +")
print("# mbarrier, MATCH = MATCH.ANY ranking, MUFU.EX2 = ex2.approx, FLO = bfind, LDS / STS = staged records; no UTMALDG /")
print("# UTCMMA / LDTM / HMMA is expected: no stage of the path is a dense contraction)")
name, counts, total = None, collections.Counter(), 0


def flush():
    if name:
        short = re.sub(r"^_ZN4pdgs\d+_GLOBAL__N__[0-9a-f_]+pdgs_\w+?_cu_[0-9a-f]+\d\d", "", name)
        body = " ".join("%s=%d" % (k, counts[k]) for k in sorted(counts) if counts[k])
        print("%-64s %5d instr  %s" % (short[:64], total, body))


for line in out.splitlines():
    m = re.match(r"\s+Function : (\S+)", line)
    if m:
        flush()
        name, counts, total = m.group(1), collections.Counter(), 0
        continue
    m = re.match(r"\s+/\*[0-9a-f]{4}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
    if m and name:
        total += 1
        op = m.group(1)
        for k in SHOW:
            if op == k or op.startswith(k + ".") or (k.count(".") and op.startswith(k)):
                counts[k] += 1
                break
flush()
