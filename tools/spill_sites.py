#!/usr/bin/env python
"""Where the production render kernel's local-memory (spill) instructions sit, by source line (no GPU needed):
   python tools/spill_sites.py > profiles/r2_spill_sites.txt
nvcc -cubin with -lineinfo, nvdisasm -g, STL / LDL counted per line of the FAST, no-record, no-counter instantiation."""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from python_raytracer_b200 import build

cubin = "/tmp/vt_spill_sites.cubin"
flags = [f for f in build.NVCC_FLAGS if f not in ("-shared", "-Xcompiler", "-fPIC", "-lz")]
subprocess.run([build._nvcc()] + flags + ["-cubin", "-o", cubin, os.path.join(build.CSRC, "voxtrace.cu")], check=True, capture_output=True)
sass = subprocess.run(["nvdisasm", "-g", cubin], capture_output=True, text=True).stdout.split("\n")
start = [i for i, l in enumerate(sass) if l.startswith(".text._ZN2vt17render_persistentILb1ELb0ELb0ELb0E")][0]
end = [i for i, l in enumerate(sass) if l.startswith(".text.") and i > start][0]
src = open(os.path.join(build.CSRC, "vt_march.cuh")).read().split("\n")
marks = [(i, m.group(1)) for i, l in enumerate(src, 1) for m in [re.match(r"__device__ .*?(\w+)\(|__global__ .*?(\w+)\(", l)] if m and m.group(1)]
marks += [(i, "render_persistent") for i, l in enumerate(src, 1) if "render_persistent(const __grid_constant__" in l]
marks.sort()


def fn(line):
    name = "?"
    for i, n in marks:
        if i <= line + 1:
            name = n
    return name


cur, st, ld, total = None, collections.Counter(), collections.Counter(), 0
for l in sass[start:end]:
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m:
        cur = (os.path.basename(m.group(1)), int(m.group(2)))
        continue
    if re.match(r"\s+/\*[0-9a-f]{4,5}\*/", l):
        total += 1
        key = fn(cur[1]) if cur[0] == "vt_march.cuh" else cur[0]
        st[key] += " STL" in l
        ld[key] += " LDL" in l
print("# render_persistent<FAST, no record, early exit, no counters>: %d instructions, %d STL, %d LDL; kernel sources sha256 %s" % (
    total, sum(st.values()), sum(ld.values()), build.kernel_source_sha256()))
print("# per function of vt_march.cuh (inlined code is attributed to the function its line belongs to) / per other file")
hot = ("march_cell", "march_leave_chunk", "rechunk", "advance", "void_step2", "cell_at", "lookup_any", "lod_snap", "snap_rel", "chunk_slot_rel")
for k in sorted(set(st) | set(ld), key=lambda k: -(st[k] + ld[k])):
    if st[k] + ld[k]:
        print("%-28s STL %3d  LDL %3d%s" % (k, st[k], ld[k], "   <-- per-sample code" if k in hot else ""))
print("# per-sample code (march_cell, march_leave_chunk, rechunk, advance, void_step2, cell_at, lookup_any, lod_snap): %d STL, %d LDL" % (
    sum(st[k] for k in hot), sum(ld[k] for k in hot)))
