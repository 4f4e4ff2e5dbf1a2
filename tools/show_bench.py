"""Prints the numbers of a bench.py JSON line (development aid).  Usage: python tools/show_bench.py file.log"""
import json
import sys

for l in open(sys.argv[1]):
    if not l.startswith("{"):
        continue
    d = json.loads(l)
    if "unavailable" in d:
        print(d); continue
    print("value", round(d["value"], 1), "ms", round(d["ms_per_step"], 3), "e2e", round(d["e2e"]["value"], 1), "n_gpus", d.get("n_gpus"),
          "cpu", d.get("cpu_baseline") and round(d["cpu_baseline"]["value"], 1), "frac", d.get("roofline") and round(d["roofline"]["frac"], 3))
    b = (d.get("config") or {}).get("builder")
    if b:
        print("  builder", b)
    for k, v in (d.get("extra") or {}).items():
        if k == "ploc_tree":
            for kk, vv in v.items():
                print("  ploc tree", kk, round(vv["value"], 1), "vs headline tree", round(vv.get("vs_headline_tree") or 0, 3), "build_ms", vv["builder"]["build_ms"])
        elif k == "naive_gpu":
            print("  naive_gpu", {a: (round(b, 2) if isinstance(b, float) else b) for a, b in v.items() if a in ("ms_per_step", "value", "this_repo_over_naive_gpu")})
        else:
            print("  extra", k, round(v["value"], 1), "kernel_ms", round(v["kernel_ms"], 3), "frac", round(v["roofline"]["frac"], 3), "build_ms", v["builder"].get("build_ms"))
    for k, v in (d.get("strong") or {}).items():
        print("  strong", k, v["assembly"], round(v["value"], 1), "ms", round(v["ms_per_step"], 3), "eff", round(v["efficiency"], 3), "verified", v["assembled_frame_verified"])
