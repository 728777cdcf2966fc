#!/bin/bash
# Bounds-checked run of the GPU suite (compute-sanitizer is not available on the GPU pool): builds libvoxtrace with
# -DVT_CHECKED=1 -- every table index of the render kernel is tested against the table size, a violation makes
# vt_render fail with the site code -- and runs the GPU tests against it.     bash tests/check_bounds.sh
set -e
cd "$(dirname "$0")/.."
python tools/build_variants.py checked:VT_CHECKED=1
export VT_LIB=$PWD/build/variants/checked.so
python -c "from python_raytracer_b200 import abi; assert abi.load().vt_build_flags() & 1, 'not the checked build'"
python -m pytest tests -m gpu -x -q -k "not bench_line"
