#!/bin/bash
# tools/ptxas_v.sh [extra nvcc flags]: builds the CUDA library (OUT=path overrides the target) and prints registers /
# spills of every trace kernel (ptxas -v), demangled.  Development only.
cd "$(dirname "$0")/.."
mkdir -p visionaray_b200/lib
nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC,-ffp-contract=off -shared -Iinclude visionaray_b200/csrc/capi.cu -o ${OUT:-visionaray_b200/lib/libvsnray_b200.so} -Xptxas -v "$@" 2>&1 | python3 -c "
import sys,re,subprocess
cur=None
for l in sys.stdin.read().split('\n'):
    m=re.search(r\"Compiling entry function '(\S+)'\",l)
    if m: cur=subprocess.run(['c++filt',m.group(1)],capture_output=True,text=True).stdout.strip(); continue
    if 'registers' in l and cur and 'trace_kernel' in cur: print(cur[:110], '|', l.split(':')[-1].strip()[:60])
    if 'spill' in l and cur and 'trace_kernel' in cur and not '0 bytes spill stores' in l: print('   ', l.strip()[:120])
    if 'error' in l: print(l)
"
