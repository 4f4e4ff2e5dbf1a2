"""Regenerates profiles/issue.json and profiles/traffic.json (what bench.py copies into roofline.issue / roofline.traffic)
from the ncu summaries of one capture round.  Usage: python tools/profiles_json.py PREFIX   (e.g. r2c: profiles/r2c_cfg2_raw.txt ...)"""
import json
import re
import sys

prefix = sys.argv[1]
KEYS = {"cfg2": "cfg2", "cfg5_closest": "cfg5", "cfg5_any": "cfg5any", "cfg3": "cfg3", "cfg4": "cfg4"}


def metrics(path):
    out = {}
    for l in open(path):
        f = re.split(r"\s{2,}", l.strip())
        if len(f) < 2 or f[0].startswith("#"):
            continue
        try:
            out[f[0]] = (float(f[-1].replace(",", "")), f[1] if len(f) == 3 else "")
        except ValueError:
            pass
    return out


def to_bytes(v, unit):
    return v * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}[unit]


issue = {"_comment": "what binds the dominant kernel, from one `ncu --set full --clock-control none` capture per workload of the same bench.py command "
                     f"(tools/gpu_profiles_c.sh; profiles/{prefix}_*_raw.txt): issue_active = smsp__issue_active.avg.pct_of_peak_sustained_active, threads_per_inst = "
                     "smsp__thread_inst_executed_per_inst_executed.ratio, useful_frac = issue_active/100 * threads_per_inst/32 (share of the SIMT issue capacity doing "
                     "useful lane work); bench.py copies the matching entry into roofline.issue"}
traffic = {"_comment": "dram__bytes_read.sum + dram__bytes_write.sum per launch of the dominant kernel, from the same captures; bench.py copies the matching entry into roofline.traffic"}
for key, name in KEYS.items():
    try:
        raw = metrics(f"profiles/{prefix}_{name}_raw.txt")
    except OSError:
        continue
    kernel = ""
    for l in open(f"profiles/{prefix}_{name}_raw.txt"):
        if "trace_kernel" in l and not kernel:
            kernel = l.strip()[:120]
    pipes = {l.split()[0]: float(l.split()[1]) for l in open(f"profiles/{prefix}_{name}_pipes.txt") if len(l.split()) == 2}
    ia = raw["smsp__issue_active.avg.pct_of_peak_sustained_active"][0]
    tpi = raw["smsp__thread_inst_executed_per_inst_executed.ratio"][0]
    dur, dunit = raw["gpu__time_duration.sum"]
    issue[key] = {"kernel": kernel, "issue_active_pct": round(ia, 2), "threads_per_inst": round(tpi, 2), "useful_frac": round(ia / 100 * tpi / 32, 3),
                  "alu_pipe_pct": round(pipes["sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active"], 1),
                  "fma_pipe_pct": round(pipes["sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active"], 1),
                  "lsu_pipe_pct": round(pipes["sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active"], 1),
                  "warps_active_pct": round(raw["sm__warps_active.avg.pct_of_peak_sustained_active"][0], 1),
                  "kernel_ms_under_ncu": round(dur * {"ms": 1, "us": 1e-3, "s": 1e3, "ns": 1e-6}.get(dunit, 1), 3),
                  "source": f"profiles/{prefix}_{name}_raw.txt"}
    rd, wr = raw["dram__bytes_read.sum"], raw["dram__bytes_write.sum"]
    traffic[key] = {"kernel": kernel, "bytes": int(to_bytes(*rd) + to_bytes(*wr)), "source": f"profiles/{prefix}_{name}_raw.txt"}
json.dump(issue, open("profiles/issue.json", "w"), indent=1)
json.dump(traffic, open("profiles/traffic.json", "w"), indent=1)
print(json.dumps({k: (v["issue_active_pct"], v["threads_per_inst"], v["alu_pipe_pct"], v["kernel_ms_under_ncu"]) for k, v in issue.items() if k != "_comment"}))
print(json.dumps({k: v["bytes"] for k, v in traffic.items() if k != "_comment"}))
