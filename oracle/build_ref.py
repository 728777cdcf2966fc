"""Recipe that makes the UNMODIFIED reference runnable where /root/reference does not exist (the GPU box), so that
bench.py can time the reference's own Python trace loop on the box's host cores (cpu_baseline kind "reference").

  python -m oracle.build_ref            (build container only; __graft_entry__.build() runs it when the tree is present)

Output goes to oracle/_ref/ only -- git-ignored, not gpurun-ignored, so it travels to the box like the built .so files --
and consists of BYTECODE compiled from the sources where they lie plus the default mod's data files:
  lib.vtbc, data.vtbc, mods/default/init.vtbc  py_compile of the reference files (pyc content; the extension is not .pyc
                                               because the gpurun snapshot leaves *.pyc behind -- refharness loads them
                                               through its own finder)
  vt_reference_init.vtbc                       the imports + class Camera + class Window of init.py (its last line opens a
                                               window, init.py:474), extracted from the AST exactly as refharness does
  assets.zip                                   the default mod's data files (config.cfg, voxels/*.gz), which data.py and the mod
                                               read at import; refharness unpacks them into a temporary directory
No reference source text is copied and nothing under oracle/_ref/ carries a reference file's name.  TEST INFRASTRUCTURE ONLY.
"""
from __future__ import annotations

import ast
import importlib.util
import os
import py_compile
import shutil
import sys

from oracle import refharness as rh

OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref")


def build(force: bool = False) -> str | None:
    src = rh.REFERENCE_ROOT
    if not rh.available():
        return None
    stamp = os.path.join(OUT, "BUILT_WITH")
    tag = "%s %s" % (sys.version.split()[0], importlib.util.MAGIC_NUMBER.hex())
    if not force and os.path.isfile(stamp) and open(stamp).read().strip() == tag:
        return OUT
    shutil.rmtree(OUT, ignore_errors=True)
    os.makedirs(os.path.join(OUT, "mods", "default"))
    for rel in ("lib.py", "data.py", os.path.join("mods", "default", "init.py")):
        py_compile.compile(os.path.join(src, rel), cfile=os.path.join(OUT, rel[:-3] + ".vtbc"), dfile=rel, doraise=True)
    with open(os.path.join(src, "init.py"), "rt") as f:
        tree = ast.parse(f.read())
    keep = [n for n in tree.body if isinstance(n, (ast.Import, ast.ImportFrom))
            or (isinstance(n, ast.ClassDef) and n.name in ("Camera", "Window"))]
    code = compile(ast.Module(body=keep, type_ignores=[]), "init.py", "exec")
    from importlib._bootstrap_external import _code_to_timestamp_pyc
    with open(os.path.join(OUT, "vt_reference_init.vtbc"), "wb") as f:
        f.write(_code_to_timestamp_pyc(code, 0, 0))
    import zipfile
    with zipfile.ZipFile(os.path.join(OUT, "assets.zip"), "w", zipfile.ZIP_DEFLATED) as z:
        z.write(os.path.join(src, "mods", "default", "config.cfg"), "mods/default/config.cfg")
        vox = os.path.join(src, "mods", "default", "voxels")
        for name in sorted(os.listdir(vox)):
            z.write(os.path.join(vox, name), "mods/default/voxels/" + name)
    for root, _, files in os.walk(OUT):
        for name in files:
            os.chmod(os.path.join(root, name), 0o644)
    with open(stamp, "w") as f:
        f.write(tag + "\n")
    return OUT


if __name__ == "__main__":
    print(build(force=True))
