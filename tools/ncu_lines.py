#!/usr/bin/env python
"""Per-source-line summary of an ncu report (needs -lineinfo and --import-source on):
   python tools/ncu_lines.py report.ncu-rep [top_n]
Prints, per CUDA source line: share of warp instructions, average active threads per instruction,
share of stall samples and the dominant stall reasons."""
import csv, io, subprocess, sys

def main(path, top=40):
    out = subprocess.run(["ncu", "-i", path, "--page", "source", "--print-source", "cuda,sass", "--csv"],
                         stdout=subprocess.PIPE, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    fname, h, lines = "", None, []
    for r in rows:
        if len(r) == 2 and r[0] == "File Path":
            fname = r[1].split("/")[-1]; continue
        if len(r) > 5 and r[0] == "Line No":
            h = r; continue
        if h and len(r) == len(h) and r[0] not in ("", "Line No") and r[2] == "-":
            lines.append((fname, r))
    ix = {n: h.index(n) for n in ["# Samples", "Instructions Executed", "Thread Instructions Executed"]}
    stall_cols = [(i, n) for i, n in enumerate(h) if n.startswith("stall_") and "Not Issued" not in n]
    f = lambda v: float(v) if v not in ("", "-") else 0.0
    tot_i = sum(f(r[ix["Instructions Executed"]]) for _, r in lines) or 1
    tot_t = sum(f(r[ix["Thread Instructions Executed"]]) for _, r in lines)
    tot_s = sum(f(r[ix["# Samples"]]) for _, r in lines) or 1
    print(f"# {path}: warp instructions {tot_i:.3g}, avg active threads {tot_t / tot_i:.2f}, stall samples {tot_s:.0f}")
    lines.sort(key=lambda x: -f(x[1][ix["# Samples"]]))
    for fn, r in lines[:top]:
        ie, te, sm = f(r[ix["Instructions Executed"]]), f(r[ix["Thread Instructions Executed"]]), f(r[ix["# Samples"]])
        st = sorted(((f(r[i]), n[6:]) for i, n in stall_cols), reverse=True)[:3]
        sts = " ".join(f"{n}:{v / max(sm, 1):.2f}" for v, n in st if v > 0)
        print(f"{fn[:20]:20s}:{r[0]:>4s} samp%={100 * sm / tot_s:5.1f} inst%={100 * ie / tot_i:5.1f} thr={te / ie if ie else 0:5.1f} [{sts}] | {r[1].strip()[:80]}")

if __name__ == "__main__":
    main(sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 40)
