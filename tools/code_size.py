#!/usr/bin/env python
"""Register / spill / code-size report of every kernel of libvoxtrace (no GPU needed):
   python tools/code_size.py > profiles/r2_code_size.txt
ptxas -v per entry point (registers, stack frame, spill bytes, shared memory) and the SASS instruction count / byte size of
each function from cuobjdump -sass (16 bytes per instruction on sm_100a)."""
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from python_raytracer_b200 import build

cubin = "/tmp/vt_code_size.cubin"
cmd = [build._nvcc()] + [f for f in build.NVCC_FLAGS if f not in ("-shared", "-Xcompiler", "-fPIC", "-lz")] + \
      ["-Xptxas", "-v", "-cubin", "-o", cubin, os.path.join(build.CSRC, "voxtrace.cu")]
res = subprocess.run(cmd, capture_output=True, text=True)
if res.returncode != 0:
    sys.exit(res.stderr)
print("# " + " ".join(cmd[1:]))
print("# kernel sources sha256 " + build.kernel_source_sha256())
name = None
for ln in res.stderr.splitlines():
    m = re.search(r"Compiling entry function '(\S+)'", ln)
    if m:
        name = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
    m = re.search(r"(\d+) bytes stack frame, (\d+) bytes spill stores, (\d+) bytes spill loads", ln)
    if m and name:
        stack = m.groups()
    m = re.search(r"Used (\d+) registers(.*)", ln)
    if m and name:
        smem = re.search(r"(\d+) bytes smem", ln)
        print("%-110s regs %3s  stack %4s B  spill st/ld %4s/%4s B  smem %5s B" % (name[:110], m.group(1), stack[0], stack[1], stack[2], smem.group(1) if smem else "0"))
        name = None
sass = subprocess.run(["cuobjdump", "-sass", cubin], capture_output=True, text=True).stdout
counts, cur = {}, None
for ln in sass.splitlines():
    m = re.search(r"Function : (\S+)", ln)
    if m:
        cur = m.group(1)
        counts[cur] = 0
    elif cur and re.match(r"\s+/\*[0-9a-f]{4,6}\*/", ln):
        counts[cur] += 1
print("\n# SASS size per function")
for k, v in sorted(counts.items(), key=lambda kv: -kv[1]):
    d = subprocess.run(["c++filt", k], capture_output=True, text=True).stdout.strip()
    print("%6d instructions %7d bytes  %s" % (v, v * 16, d[:120]))
