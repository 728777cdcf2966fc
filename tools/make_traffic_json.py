#!/usr/bin/env python
"""Per-launch DRAM traffic and issue statistics of the render kernel from a tools/ncu_summary.py text -> the JSON
bench.py reads for roofline.traffic (profiles/r*_traffic.json):
   python tools/make_traffic_json.py summary.txt <name to cite> [sha256 of the kernel sources the capture was taken from]"""
import json, re, sys
txt = open(sys.argv[1]).read()
cite = sys.argv[2] if len(sys.argv) > 2 else sys.argv[1]
launches = txt.split("-- launch ")[1:]


def val(block, key):
    m = re.search(re.escape(key) + r" \[[^\]]*\] = ([-0-9.eE+,]+)", block)
    return float(m.group(1).replace(",", "")) if m else None


def to_bytes(block, key):
    m = re.search(re.escape(key) + r" \[([^\]]*)\] = ([-0-9.eE+,]+)", block)
    if not m:
        return None
    unit, v = m.group(1).lower(), float(m.group(2).replace(",", ""))
    return v * {"byte": 1, "kbyte": 1e3, "mbyte": 1e6, "gbyte": 1e9}.get(unit, 1)


def to_ms(block, key):
    m = re.search(re.escape(key) + r" \[([^\]]*)\] = ([-0-9.eE+,]+)", block)
    if not m:
        return None
    unit, v = m.group(1).lower(), float(m.group(2).replace(",", ""))
    return v * {"ns": 1e-6, "us": 1e-3, "usecond": 1e-3, "ms": 1.0, "msecond": 1.0, "s": 1e3, "second": 1e3, "nsecond": 1e-6}.get(unit, 1e-6)


rd = [to_bytes(b, "dram__bytes_read.sum") for b in launches]
wr = [to_bytes(b, "dram__bytes_write.sum") for b in launches]
out = {
    "kernel": "render_persistent<FAST>",
    **({"kernel_sha256": sys.argv[3]} if len(sys.argv) > 3 else {}),
    "source": "%s (ncu --set full --clock-control none, bench.py --steps 4 --warmup 3, the four timed launches = the four camera poses)" % cite,
    "dram_read_bytes_per_launch": rd,
    "dram_write_bytes_per_launch": wr,
    "traffic_bytes_per_launch_mean": sum(a + b for a, b in zip(rd, wr)) / max(1, len(rd)),
    "ncu_ms_per_launch": [to_ms(b, "gpu__time_duration.sum") for b in launches],
    "issue_active_pct": [val(b, "smsp__issue_active.avg.pct_of_peak_sustained_active") for b in launches],
    "lanes_per_instruction": [val(b, "smsp__thread_inst_executed_per_inst_executed.ratio") for b in launches],
    "warps_active_pct": [val(b, "sm__warps_active.avg.pct_of_peak_sustained_active") for b in launches],
    "l1_hit_pct": [val(b, "l1tex__t_sector_hit_rate.pct") for b in launches],
    "l2_hit_pct": [val(b, "lts__t_sector_hit_rate.pct") for b in launches],
    "fp64_pipe_pct": [val(b, "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active") for b in launches],
}
print(json.dumps(out, indent=1))
