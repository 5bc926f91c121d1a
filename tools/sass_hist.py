#!/usr/bin/env python
"""Opcode histogram of one kernel's SASS (whole function, and its hottest loop =
the longest backward-branch span). usage: sass_hist.py <lib.so> <mangled-substring>"""
import re, subprocess, sys, collections
lib, key = sys.argv[1], sys.argv[2]
txt = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
cur, ins = None, []
for line in txt.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        cur = m.group(1); continue
    if cur and key in cur:
        m = re.match(r"\s+/\*([0-9a-f]{4,5})\*/\s+(.*?);", line)
        if m:
            ins.append((int(m.group(1), 16), m.group(2)))
def op(s):
    s = re.sub(r"^@!?U?P\d+\s+", "", s)
    return s.split()[0]
# loop: backward BRA with the largest span
best = (0, 0, 0)
for a, s in ins:
    m = re.search(r"BRA\S*\s+.*0x([0-9a-f]+)", s)
    if m and op(s).startswith("BRA"):
        t = int(m.group(1), 16)
        if t < a and a - t > best[0]:
            best = (a - t, t, a)
_, lo, hi = best
body = [s for a, s in ins if lo <= a <= hi]
def hist(lst):
    c = collections.Counter(re.sub(r"\..*", "", op(s)) for s in lst)
    return c
print(f"function insts {len(ins)}, loop [{lo:#x},{hi:#x}] insts {len(body)}")
h = hist(body)
pipes = {"fma": ["FFMA2", "FFMA", "FMUL", "FADD", "FMUL2", "FADD2", "IMAD", "HFMA2"],
         "alu": ["FMNMX", "FMNMX3", "FSEL", "FSETP", "PRMT", "LOP3", "IADD3", "ISETP", "SEL", "SHF", "MOV", "IADD", "LEA", "PLOP3", "VIADD", "I2FP", "F2FP", "FSET", "IABS", "VIMNMX", "IMNMX", "UMOV"],
         "xu": ["MUFU", "F2I", "I2F", "F2F"], "lsu": ["LDS", "STS", "LDG", "STG", "LDGSTS", "LDL", "STL", "LDC", "LDGDEPBAR", "DEPBAR"]}
for k, v in sorted(h.items(), key=lambda kv: -kv[1]):
    pp = next((p for p, l in pipes.items() if k in l), "?")
    print(f"{v:5d} {k:10s} {pp}")
