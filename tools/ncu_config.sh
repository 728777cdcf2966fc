#!/bin/bash
# ncu --set full capture of one render launch of a named workload (c3 / c4 / c5) on the GPU box:
#   bash tools/ncu_config.sh c4 [tag]   -> gpurun_out/<tag>_<wl>_ncu.txt (per-launch digest) and <tag>_<wl>_hot_lines.txt
set -e
cd "$(dirname "$0")/.."
WL=${1:-c4}; TAG=${2:-r2}
OUT=gpurun_out; mkdir -p $OUT
ncu --set full --clock-control none --import-source on --kernel-name regex:render_persistent --launch-skip 3 --launch-count 1 \
    -f -o $OUT/${TAG}_${WL} python tools/cbench.py WL=$WL > $OUT/${TAG}_${WL}_ncu_run.log 2>&1
ncu -i $OUT/${TAG}_${WL}.ncu-rep --page raw --csv | python tools/ncu_summary.py "render_persistent, tools/cbench.py WL=$WL, one launch (first pose)" > $OUT/${TAG}_${WL}_ncu.txt
ncu -i $OUT/${TAG}_${WL}.ncu-rep --page source --print-source cuda,sass --csv > $OUT/${TAG}_${WL}_source.csv 2>/dev/null || true
python tools/ncu_by_line.py $OUT/${TAG}_${WL}_source.csv 40 > $OUT/${TAG}_${WL}_hot_lines.txt 2>/dev/null || true
rm -f $OUT/${TAG}_${WL}_source.csv $OUT/${TAG}_${WL}.ncu-rep
cat $OUT/${TAG}_${WL}_ncu.txt
