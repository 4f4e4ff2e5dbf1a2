#!/usr/bin/env python
"""Executed warp instructions per SASS opcode of an ncu report (source page), with the average active
threads: which opcodes load the alu / fma pipes.   python tools/ncu_opcodes.py report.ncu-rep [top_n]"""
import collections, csv, io, subprocess, sys


def main(path, top=40):
    out = subprocess.run(["ncu", "-i", path, "--page", "source", "--print-source", "sass", "--csv"], stdout=subprocess.PIPE, text=True).stdout
    h = None
    acc = collections.defaultdict(lambda: [0.0, 0.0])
    for r in csv.reader(io.StringIO(out)):
        if len(r) > 5 and "Source" in r and "Instructions Executed" in r:
            h = r; si = h.index("Source"); ii = h.index("Instructions Executed"); ti = h.index("Thread Instructions Executed"); continue
        if not h or len(r) != len(h):
            continue
        try:
            ie, te = float(r[ii]), float(r[ti])
        except ValueError:
            continue
        toks = r[si].split()
        if not toks:
            continue
        op = toks[1] if toks[0].startswith("@") and len(toks) > 1 else toks[0]
        a = acc[op.rstrip(";").split(".")[0]]; a[0] += ie; a[1] += te
    tot = sum(a[0] for a in acc.values()) or 1
    print(f"# {path}: {tot:.4g} warp instructions")
    for k, a in sorted(acc.items(), key=lambda x: -x[1][0])[:top]:
        print(f"  {k:12s} inst%={100 * a[0] / tot:5.1f}  active_threads={a[1] / a[0] if a[0] else 0:5.1f}")


if __name__ == "__main__":
    main(sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 40)
