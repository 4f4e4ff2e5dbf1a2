#!/usr/bin/env python
"""Instruction / stall-sample shares per code region of the trace kernel (file + line ranges)."""
import csv, io, subprocess, sys
def load(path):
    """-> header, [(file, line, row-of-the-SASS-instruction)]: one entry per (SASS address, source line it is listed under).
    An inlined instruction is listed under EVERY frame of its inline stack; dedupe() keeps one region per address."""
    out = subprocess.run(["ncu","-i",path,"--page","source","--print-source","cuda,sass","--csv"],stdout=subprocess.PIPE,text=True).stdout
    rows=list(csv.reader(io.StringIO(out))); fname="";h=None;L=[];cur=None
    for r in rows:
        if len(r)==2 and r[0]=="File Path": fname=r[1].split("/")[-1]; continue
        if len(r)>5 and r[0]=="Line No": h=r; continue
        if not h or len(r)!=len(h): continue
        if r[0] not in("","Line No") and r[2]=="-": cur=(fname,int(r[0]),r[1]); continue
        if r[0]=="" and r[2].startswith("0x") and cur: L.append((cur[0],cur[1],r))
    return h,L
# region priority: the most specific frame of an inlined instruction wins
_PRIO=["node_test","isect_triangle / rn arithmetic (vsb_common.cuh)","prim_test (record decode, accept)","hit record of the winner (once per ray)","record loads","stack","ray_init","shade programs"]
def dedupe(L):
    """one (region, row) per SASS address"""
    by={}
    for fn,ln,r in L:
        reg=region(fn,ln,"")
        a=by.get(r[2])
        if a is None: by[r[2]]=[reg,r]; continue
        pa=_PRIO.index(a[0]) if a[0] in _PRIO else 99; pb=_PRIO.index(reg) if reg in _PRIO else 99
        if pb<pa: a[0]=reg
    return list(by.values())
import os, re
def _function_starts():
    """line number of every top-level function of bvh8_traverse.cuh (regions follow the source, not fixed lines)"""
    path=os.path.join(os.path.dirname(os.path.abspath(__file__)),"..","visionaray_b200","csrc","bvh8_traverse.cuh")
    out=[]
    for i,l in enumerate(open(path),1):
        m=re.match(r"^(?:VSB_HD|__device__ __forceinline__)\s+[\w: ]+?\s+(\w+)\(",l)
        if m: out.append((i,m.group(1)))
    return out
_FN=_function_starts()
_KIND={"test_node":"node_test","perm_xor8":"node_test","test_node_slots":"node_test","test_node_lut":"node_test","expand_nibbles":"node_test","funnel_shift_in_sign":"node_test","perm_lut_fill":"kernel loop (service, refill, program)","prim_record":"queue: append + drain decision","test_prim":"prim_test (record decode, accept)","prim_candidate":"prim_test (record decode, accept)","hit_from_record":"hit record of the winner (once per ray)","load_record":"record loads","ldg_hint":"record loads",
       "ray_ctx_init":"ray_init","stack_push":"stack","stack_pop":"stack","trav_begin":"stack","trav_step":"trav_step control"}
def _kernel_markers():
    """regions of trace_kernel_q (kernels.cuh) by the marker comments of its loop"""
    path=os.path.join(os.path.dirname(os.path.abspath(__file__)),"..","visionaray_b200","csrc","kernels.cuh")
    marks=[("auto drain = ","queue: drain (pair fetch, ray from shared, key merge)"),("// ---- continuations of finished rays","kernel loop (service, refill, program)"),
           ("// ---- node phase","queue: node phase control"),("// ---- append the selected","queue: append + drain decision"),("// ---- append one GROUP","queue: append + drain decision"),
           ("if (drained && (state","queue: limit refresh / state"),("if (tl.ray_count)","kernel loop (service, refill, program)")]
    out=[]; inq=False
    for i,l in enumerate(open(path),1):
        if "trace_kernel_q(const TraceLaunch" in l: inq=True
        if "struct RayListProgram" in l: inq=False
        if not inq: continue
        for key,name in marks:
            if key in l: out.append((i,name))
    return out
_KM=_kernel_markers()
def _kernel_region(ln):
    name="kernel loop (service, refill, program)"
    for start,n in _KM:
        if start<=ln: name=n
    return name
def region(fn, ln, src):
    if fn=="bvh8_traverse.cuh":
        name=""
        for start,n in _FN:
            if start<=ln: name=n
        return _KIND.get(name,"trav_step control")
    if fn in ("bvh8.cuh","cuda_fp16.hpp"): return "node_test"      # cuda_fp16.hpp: the half2 -> float conversions of the plane bytes
    if fn=="vsb_common.cuh": return "isect_triangle / rn arithmetic (vsb_common.cuh)"
    if fn=="kernels.cuh": return _kernel_region(ln)
    if fn=="shade.cuh": return "shade programs"
    return fn
if __name__=="__main__":
    path=sys.argv[1]
    h,L=load(path)
    ii=h.index("Instructions Executed"); ti=h.index("Thread Instructions Executed"); si=h.index("# Samples")
    f=lambda v: float(v) if v not in("","-") else 0.0
    g={}
    for reg,r in dedupe(L):
        a=g.setdefault(reg,[0,0,0]); a[0]+=f(r[ii]); a[1]+=f(r[ti]); a[2]+=f(r[si])
    T=sum(a[0] for a in g.values()); S=sum(a[2] for a in g.values())
    print(f"# {path}: warp instructions {T:.4g} (every SASS address counted once; inlined code goes to its most specific region)")
    for k,a in sorted(g.items(), key=lambda x:-x[1][0]): print(f"  {k:40s} inst%={100*a[0]/T:5.1f} active_threads={a[1]/a[0] if a[0] else 0:5.1f} stall_samples%={100*a[2]/S:5.1f}")
