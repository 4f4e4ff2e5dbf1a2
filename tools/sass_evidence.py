#!/usr/bin/env python
"""SASS evidence for DESIGN.md 4.1: per trace kernel, the instruction count, the 128-bit record loads (LDG.E.128), the
local-memory instructions (LDL / STL) inside and outside the main loop (= the widest backward branch), shared-memory
atomics, and an excerpt around the node record's loads.   python tools/sass_evidence.py [lib.so] > profiles/..."""
import re, subprocess, sys

lib = sys.argv[1] if len(sys.argv) > 1 else "visionaray_b200/lib/libvsnray_b200.so"
out = subprocess.run(["cuobjdump", "-sass", lib], stdout=subprocess.PIPE, text=True).stdout
funcs, cur = {}, None
for l in out.split("\n"):
    m = re.search(r"Function : (\S+)", l)
    if m:
        cur = m.group(1); funcs[cur] = []; continue
    m = re.match(r"\s+/\*([0-9a-f]{4,6})\*/\s+(.*?);", l)
    if m and cur:
        funcs[cur].append((int(m.group(1), 16), m.group(2).strip()))
want = [("trace_kernel_q<FrameProgram<1,0,0>> (cfg2: Whitted, lean instantiation)", "trace_kernel_qINS_12FrameProgramILi1ELb0ELb0E"),
        ("trace_kernel_q<FrameProgram<3,0,0>> (cfg4: path tracer)", "trace_kernel_qINS_12FrameProgramILi3ELb0ELb0E"),
        ("trace_kernel_q<FrameProgram<2,0,0>> (cfg3: AO)", "trace_kernel_qINS_12FrameProgramILi2ELb0ELb0E"),
        ("trace_kernel_q<RayListProgram> (cfg5)", "trace_kernel_qINS_14RayListProgram")]
print(f"# cuobjdump -sass {lib} (sm_100a), summarised by tools/sass_evidence.py")
for title, key in want:
    name = next((n for n in funcs if key in n), None)
    if not name:
        print(f"\n## {title}: not found"); continue
    ins = funcs[name]
    loop = (0, 0)
    for a, t in ins:
        m = re.search(r"\bBRA(?:\.\w+)*\s+(?:\w+,\s*)?0x([0-9a-f]+)", t)
        if m and int(m.group(1), 16) < a and a - int(m.group(1), 16) > loop[1] - loop[0]:
            loop = (int(m.group(1), 16), a)
    inside = lambda a: loop[0] <= a <= loop[1]
    cnt = lambda pat, f=lambda a: True: sum(1 for a, t in ins if re.search(pat, t) and f(a))
    print(f"\n## {title}\n{name}")
    print(f"instructions {len(ins)}; main loop 0x{loop[0]:x}..0x{loop[1]:x} = {sum(1 for a, _ in ins if inside(a))} instructions")
    print(f"LDG.E.128 (record loads: 5 per node, 3 per triangle) {cnt(r'LDG\.E\.128')} | other LDG {cnt(r'LDG\.E(?!\.128)')}")
    print(f"local memory inside the loop: LDL {cnt(r'^(@!?U?P\d+ )?LDL', inside)} STL {cnt(r'^(@!?U?P\d+ )?STL', inside)}   outside: LDL {cnt(r'^(@!?U?P\d+ )?LDL', lambda a: not inside(a))} STL {cnt(r'^(@!?U?P\d+ )?STL', lambda a: not inside(a))}")
    print(f"shared memory: LDS {cnt(r'LDS')} STS {cnt(r'STS')} ATOMS {cnt(r'ATOMS')} | PRMT {cnt(r'PRMT')} HADD2.F32 {cnt(r'HADD2\.F32')} FFMA {cnt(r'FFMA')} FMNMX3 {cnt(r'FMNMX3')}")
    for a, t in ins:
        if re.search(r"^(@!?U?P\d+ )?(LDL|STL)", t) and inside(a):
            print(f"    0x{a:04x}  {t}")
    i0 = next((i for i, (a, t) in enumerate(ins) if "LDG.E.128" in t and "+0x40" in t), None)
    if i0 is not None:
        print("  excerpt: the five 128-bit loads of a node record and the start of the node test")
        for a, t in ins[max(i0 - 6, 0):i0 + 34]:
            print(f"    0x{a:04x}  {t}")
