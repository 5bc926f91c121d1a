#!/usr/bin/env python
"""Register-file read traffic of a kernel's hot loop: 32-bit register source
operands per loop iteration, reuse-cache hits subtracted (an operand is a hit when
the previous instruction carried the same register in the same source slot with
the .reuse flag). usage: sass_rf.py <lib.so> <mangled-substring> [--steps N]"""
import argparse, re, subprocess, collections
from sass_hist import load, opcode, base

def srcs(s):
    body = re.sub(r"^@!?U?P\d+\s+", "", s)
    parts = body.split(None, 1)
    if len(parts) < 2: return []
    ops = [o.strip() for o in parts[1].split(",")]
    op = base(parts[0])
    # first operand is the destination for everything but stores / branches
    if op in ("STG", "STS", "STL", "LDGSTS", "BRA", "EXIT", "DEPBAR", "LDGDEPBAR", "BAR", "RED", "ATOMG"):
        s_ops = ops
    else:
        s_ops = ops[1:]
        if op in ("FSETP", "ISETP", "PLOP3", "UISETP") and len(s_ops) > 1:
            s_ops = s_ops[1:]          # second predicate destination
    out = []
    for slot, o in enumerate(s_ops):
        for m in re.finditer(r"(?<![UP\w])R(\d+)(\.reuse)?(\.64|\.F32x2\S*)?", o):
            if m.group(1) == "Z": continue
            n = 2 if (m.group(3) or ("desc" in o and ".64" in o)) else 1
            out.append((slot, int(m.group(1)), n, bool(m.group(2))))
    return out

ap = argparse.ArgumentParser(); ap.add_argument("lib"); ap.add_argument("key"); ap.add_argument("--steps", type=int, default=1)
a = ap.parse_args()
name, ins = load(a.lib, a.key)
best = (0, 0, 0)
for addr, s in ins:
    m = re.search(r"BRA\S*\s+.*0x([0-9a-f]+)", s)
    if m and opcode(s).startswith("BRA"):
        t = int(m.group(1), 16)
        if t < addr and addr - t > best[0]: best = (addr - t, t, addr)
_, lo, hi = best
body = [s for addr, s in ins if lo <= addr <= hi]
total = hits = 0
prev = {}
per_op = collections.Counter()
for s in body:
    cur = {}
    for slot, r, n, reuse in srcs(s):
        if prev.get(slot) == r:
            hits += n
        else:
            total += n; per_op[base(opcode(s))] += n
        if reuse: cur[slot] = r
    prev = cur
print(f"{name}: loop {len(body)} instructions, {total} register reads (+{hits} reuse hits) per iteration;"
      f" per row step: {total / a.steps:.0f} reads = {total / a.steps / 2:.0f} clk at 2 reads/clk, {len(body) / a.steps:.0f} issue slots")
print({k: v for k, v in per_op.most_common(8)})
