#!/bin/bash
# Round-end evidence of the render kernel, run on the GPU box (gpurun): plain bench line, launch list, one
# `ncu --set full` capture of four timed launches (the four camera poses), per-launch summary, hot source lines and the
# DRAM traffic file bench.py reads.  Outputs under gpurun_out/; copy what is to be judged into profiles/.
#   bash tools/final_capture.sh [tag] [round]      (the traffic file carries the sha256 of the kernel sources it was taken from)
set -e
cd "$(dirname "$0")/.."
TAG=${1:-final}
ROUND=${2:-r2}
SHA=$(python -c "from python_raytracer_b200 import build; print(build.kernel_source_sha256())")
OUT=gpurun_out
mkdir -p $OUT
python bench.py > $OUT/${TAG}_bench_n1.json 2> $OUT/${TAG}_bench_n1.err
python bench.py --steps 4 --warmup 3 --no-configs --no-cpu-baseline > $OUT/${TAG}_bench_short.json 2>/dev/null
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/${TAG}_launches.csv \
    python bench.py --steps 20 --warmup 3 --no-configs --no-cpu-baseline > $OUT/${TAG}_launches_run.log 2>&1
ncu --set full --clock-control none --import-source on --kernel-name regex:render_persistent --launch-skip 3 --launch-count 4 \
    -f -o $OUT/${TAG}_render python bench.py --steps 4 --warmup 3 --no-configs --no-cpu-baseline > $OUT/${TAG}_ncu_run.log 2>&1
ncu -i $OUT/${TAG}_render.ncu-rep --page raw --csv | python tools/ncu_summary.py "render_persistent, bench.py --steps 4 --warmup 3, four timed launches" \
    > $OUT/${TAG}_render_persistent_ncu.txt
ncu -i $OUT/${TAG}_render.ncu-rep --page source --print-source cuda,sass --csv > $OUT/${TAG}_source.csv 2>/dev/null || true
python tools/ncu_by_line.py $OUT/${TAG}_source.csv 40 > $OUT/${TAG}_hot_lines.txt 2>/dev/null || true
python tools/make_traffic_json.py $OUT/${TAG}_render_persistent_ncu.txt "profiles/${ROUND}_render_persistent_ncu.txt" $SHA > $OUT/${TAG}_traffic.json
rm -f $OUT/${TAG}_source.csv
ls -la $OUT | tail -12
