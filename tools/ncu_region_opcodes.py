#!/usr/bin/env python
"""Opcode histogram of ONE code region of the trace kernel (regions as in tools/ncu_groups.py), with the pipe each
opcode issues to (alu / fma / other): which instructions to attack when one pipe is the limiter.
Usage: python tools/ncu_region_opcodes.py report.ncu-rep [region-substring]"""
import re
import sys

argv = sys.argv
import importlib.util, os
spec = importlib.util.spec_from_file_location("ncu_groups", os.path.join(os.path.dirname(os.path.abspath(__file__)), "ncu_groups.py"))
G = importlib.util.module_from_spec(spec)
spec.loader.exec_module(G)
want = argv[2] if len(argv) > 2 else "node_test"
ALU = {"LOP3", "SHF", "PRMT", "FMNMX", "FMNMX3", "ISETP", "SEL", "IADD3", "BMSK", "FSETP", "LEA", "VIADD", "MOV", "POPC", "FLO", "PLOP3", "VIMNMX", "IABS", "FSEL", "SGXT", "R2P", "P2R", "BREV", "VIMNMX3", "HMNMX2", "HSETP2", "IMNMX", "LOP", "FCHK", "FSET"}
FMA = {"FFMA", "FMUL", "FADD", "IMAD", "HADD2", "HFMA2", "HMUL2"}
h, L = G.load(argv[1])
ii = h.index("Instructions Executed"); ti = h.index("Thread Instructions Executed")
f = lambda v: float(v) if v not in ("", "-") else 0.0
ops = {}
total = 0.0
for reg, r in G.dedupe(L):
    n = f(r[ii]); total += n
    if want not in reg:
        continue
    m = re.match(r"\s*(?:@!?U?P\d+\s+)?([A-Z0-9_]+)", r[3])
    op = m.group(1) if m else "?"
    a = ops.setdefault(op, [0.0, 0.0]); a[0] += n; a[1] += f(r[ti])
T = sum(a[0] for a in ops.values())
print(f"# region '{want}': {100 * T / total:.1f} % of {total:.4g} warp instructions")
pipes = {"alu": 0.0, "fma": 0.0, "other": 0.0}
for op, a in sorted(ops.items(), key=lambda x: -x[1][0]):
    pipe = "alu" if op in ALU else ("fma" if op in FMA else "other")
    pipes[pipe] += a[0]
    if a[0] / T >= 0.004:
        print(f"  {op:10s} {pipe:5s} {100 * a[0] / T:5.1f} % of region  {100 * a[0] / total:5.2f} % of kernel  lanes {a[1] / a[0]:5.1f}")
print("  pipes:", {k: f"{100 * v / T:.1f} % of region, {100 * v / total:.1f} % of kernel" for k, v in pipes.items()})
