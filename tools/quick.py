#!/usr/bin/env python
"""Print the few numbers of a bench.py JSON line (stdin) on one line."""
import json, sys
tag = sys.argv[1] if len(sys.argv) > 1 else ""
for l in sys.stdin:
    l = l.strip()
    if l.startswith("{"):
        d = json.loads(l)
        r = d.get("roofline") or {}
        print(f"{tag:12s} value={d['value']:9.1f} ms/step={d['ms_per_step']:8.3f} kernel_ms={r.get('kernel_ms', 0):8.3f} frac={r.get('frac', 0):.3f} e2e={d['e2e']['value']:.1f} e2e_ms={d['e2e'].get('ms_per_step', 0):.2f} e2e_trace_ms={d['e2e'].get('trace_ms', 0):.2f} clk={(d.get('clocks') or {}).get('sm_mhz')}/{(d.get('clocks') or {}).get('sm_max_mhz')} {(d.get('clocks') or {}).get('reasons')}")
