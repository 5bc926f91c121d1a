#!/usr/bin/env python
"""Opcode histogram and pipe-time model of one kernel's hot loop (the longest
backward-branch span of its SASS).

    sass_hist.py <lib.so> <mangled-substring> [--steps N] [--pixels P] [--ms T --frames F --w W --h H]

Per loop iteration it prints the instruction count per opcode and pipe and the
clocks each pipe of an SM sub-partition is busy, from the costs measured on the B200
with tests/cuda/pipebench.cu (profiles/r1_pipebench_operands_b200.txt,
profiles/r2_pipebench_rf_b200.txt):

    FFMA2 / FADD2 / FMUL2   max(2.07, reads / 2) clk, reads = 32-bit register words
                            the instruction fetches (2 per 64-bit pair source, 1 per
                            scalar source; immediates, uniform registers, RZ and a
                            repeated register cost nothing): two pairs 2.07, pair +
                            pair + scalar 2.5 (3.5 measured back to back), three
                            pairs 3.0 (3.13 measured)
    FFMA / FADD / FMUL      1.04 clk (1.5 with three distinct register sources)
    IMAD / IMAD.WIDE        1.04 clk (FMA pipe)
    ALU pipe (FSEL, FSETP, FMNMX, PRMT, LOP3, IADD3, MOV, ISETP, ...)   2 clk
    MUFU                    8 clk (XU pipe)
    issue                   1 clk per instruction of any kind

--steps N: row steps per loop iteration (loops instantiated N times with renamed
registers); --pixels P: pixel pairs a row step produces per thread (4 columns: 4).
The largest of the per-pipe times is the kernel's instruction-bound floor; with
--ms it is compared with a measured launch."""
import argparse
import collections
import re
import subprocess

PIPES = {
    "fma": ["FFMA2", "FFMA", "FMUL", "FADD", "FMUL2", "FADD2", "IMAD", "HFMA2"],
    "alu": ["FMNMX", "FMNMX3", "FSEL", "FSETP", "PRMT", "LOP3", "IADD3", "ISETP", "SEL", "SHF", "MOV",
            "IADD", "LEA", "PLOP3", "VIADD", "I2FP", "F2FP", "FSET", "IABS", "VIMNMX", "VIMNMX3", "IMNMX",
            "CS2R", "P2R", "R2P", "FCHK", "LOP", "SGXT", "BMSK", "POPC", "FLO"],
    "xu": ["MUFU", "F2I", "I2F", "F2F"],
    "lsu": ["LDS", "STS", "LDG", "STG", "LDGSTS", "LDL", "STL", "LDC", "LDGDEPBAR", "DEPBAR", "ATOMG", "RED"],
    "uniform": ["UMOV", "UIADD3", "ULDC", "UISETP", "ULOP3", "USEL", "UIMAD", "USHF", "ULEA", "UPRMT", "LDCU",
                "S2UR", "R2UR", "UFLO", "UPLOP3"],
}


def pipe_of(op):
    return next((p for p, l in PIPES.items() if op in l), "other")


def load(lib, key):
    txt = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
    cur, ins, name = None, [], None
    for line in txt.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            cur = m.group(1)
            continue
        if cur and key in cur:
            name = cur
            m = re.match(r"\s+/\*([0-9a-f]{4,5})\*/\s+(.*?);", line)
            if m:
                ins.append((int(m.group(1), 16), m.group(2)))
    return name, ins


def opcode(s):
    s = re.sub(r"^@!?U?P\d+\s+", "", s)
    return s.split()[0]


def base(op):
    return re.sub(r"\..*", "", op)


def packed_cost(s):
    """clocks of a packed FP32 instruction from its source operand form"""
    ops = s.split(None, 1)[1] if " " in s else ""
    srcs = [o.strip() for o in ops.split(",")][1:]
    regs = set()
    for o in srcs:
        m = re.match(r"[-|]*R(\d+)(\.reuse)?(\.F32x2\S*|\.F32)?", o)
        if m and not o.startswith("UR"):
            regs.add((m.group(1), bool(m.group(3) and "x2" in m.group(3)) or m.group(3) is None))
    pairs = sum(1 for _, wide in regs if wide)
    scalars = len(regs) - pairs
    return max(2.07, 0.5 * (2 * pairs + scalars)), "P" * pairs + "s" * scalars


def scalar_cost(s):
    ops = s.split(None, 1)[1] if " " in s else ""
    srcs = [o.strip() for o in ops.split(",")][1:]
    regs = {re.match(r"[-|]*R(\d+)", o).group(1) for o in srcs if re.match(r"[-|]*R\d+", o)}
    return 1.5 if len(regs) >= 3 else 1.04


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("lib")
    ap.add_argument("key")
    ap.add_argument("--steps", type=int, default=1)
    ap.add_argument("--pixels", type=int, default=4, help="pixel pairs per row step and thread")
    ap.add_argument("--ms", type=float, default=0.0)
    ap.add_argument("--frames", type=int, default=0)
    ap.add_argument("--w", type=int, default=0)
    ap.add_argument("--h", type=int, default=0)
    ap.add_argument("--mhz", type=float, default=1965.0)
    a = ap.parse_args()
    name, ins = load(a.lib, a.key)
    if not ins:
        raise SystemExit(f"no function matching {a.key}")
    best = (0, 0, 0)
    for addr, s in ins:
        m = re.search(r"BRA\S*\s+.*0x([0-9a-f]+)", s)
        if m and opcode(s).startswith("BRA"):
            t = int(m.group(1), 16)
            if t < addr and addr - t > best[0]:
                best = (addr - t, t, addr)
    _, lo, hi = best
    body = [s for addr, s in ins if lo <= addr <= hi]
    demangled = subprocess.run(["c++filt", name], capture_output=True, text=True).stdout.strip()
    print(f"kernel   {demangled}")
    print(f"function {len(ins)} instructions, hot loop [{lo:#x}, {hi:#x}] {len(body)} instructions"
          f" = {a.steps} row step(s) of {a.pixels} pixel pairs per thread")
    hist = collections.Counter(base(opcode(s)) for s in body)
    clocks = collections.Counter()
    forms = collections.Counter()
    for s in body:
        op = base(opcode(s))
        p = pipe_of(op)
        if op in ("FFMA2", "FADD2", "FMUL2"):
            c, form = packed_cost(s)
            forms[f"{form or '-'} {c:.2f}"] += 1
            clocks["fma"] += c
        elif op in ("FFMA", "FADD", "FMUL"):
            clocks["fma"] += scalar_cost(s)
        elif p == "fma":
            clocks["fma"] += 1.04
        elif p == "alu":
            clocks["alu"] += 2.0
        elif p == "xu":
            clocks["xu"] += 8.0
        elif p == "lsu":
            clocks["lsu"] += 2.0
    clocks["issue"] = float(len(body))
    for k, v in sorted(hist.items(), key=lambda kv: -kv[1]):
        print(f"{v:6d} {k:10s} {pipe_of(k)}")
    print("packed FP32 forms:", dict(forms))
    per_step = {k: v / a.steps for k, v in clocks.items()}
    print("clocks per row step, per pipe of one SM sub-partition (one warp):",
          {k: round(v, 1) for k, v in per_step.items()})
    floor_clk = max(per_step.values())
    bound = max(per_step, key=per_step.get)
    packed = sum(v for k, v in hist.items() if k in ("FFMA2", "FADD2", "FMUL2")) / a.steps
    print(f"packed FP32 instructions per pixel pair: {packed / a.pixels:.1f}; "
          f"instruction-bound floor {floor_clk:.0f} clk per row step ({bound} pipe)")
    if a.ms > 0 and a.frames > 0:
        px_pairs = (a.frames / 2.0) * a.w * a.h
        warp_steps = px_pairs / (32.0 * a.pixels)            # row steps of one warp
        smsp = 148 * 4
        floor_ms = warp_steps * floor_clk / smsp / (a.mhz * 1e3)
        print(f"measured {a.ms:.4f} ms for {a.frames} frames {a.w}x{a.h}: floor {floor_ms:.4f} ms at {a.mhz:.0f} MHz "
              f"-> {100 * floor_ms / a.ms:.0f} % of the instruction-bound floor "
              f"({a.ms / warp_steps * smsp * a.mhz * 1e3:.0f} clk per row step measured)")


if __name__ == "__main__":
    main()
