#!/usr/bin/env python
"""Build tuning variants of libvoxtrace.so into build/variants/ (git-ignored, travels to the GPU box):
   python tools/build_variants.py name:DEF=VAL,DEF=VAL ...      then   python tools/kbench.py LIB=build/variants/name.so,..."""
import os, sys
from concurrent.futures import ThreadPoolExecutor
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from python_raytracer_b200 import build
os.makedirs(os.path.join(ROOT, "build", "variants"), exist_ok=True)
def one(spec):
    name, _, defs = spec.partition(":")
    out = os.path.join(ROOT, "build", "variants", name + ".so")
    build.build_library(force=True, defines=[d for d in defs.split(",") if d], out=out)
    return out
with ThreadPoolExecutor(8) as ex:
    for o in ex.map(one, sys.argv[1:]):
        print(o)
