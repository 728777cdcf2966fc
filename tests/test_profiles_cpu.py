"""The numbers bench.py quotes from profiles/ can be re-derived from the committed ncu digests (no GPU needed)."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


import pytest


@pytest.mark.parametrize("name", ["r1_traffic.json", "r2_traffic.json"])
def test_traffic_json_follows_from_the_committed_ncu_summary(name):
    tj = json.load(open(os.path.join(ROOT, "profiles", name)))
    cited = tj["source"].split(" ")[0]
    assert os.path.exists(os.path.join(ROOT, cited)), cited
    extra = [tj["kernel_sha256"]] if "kernel_sha256" in tj else []
    out = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "make_traffic_json.py"), os.path.join(ROOT, cited), cited] + extra,
                         capture_output=True, text=True, check=True).stdout
    again = json.loads(out)
    assert again == tj
    n = len(tj["dram_read_bytes_per_launch"])
    assert n == 4                                               # the four camera poses of the bench workload
    mean = sum(a + b for a, b in zip(tj["dram_read_bytes_per_launch"], tj["dram_write_bytes_per_launch"])) / n
    assert abs(mean - tj["traffic_bytes_per_launch_mean"]) < 1.0


def test_scene_row_roofline_file_is_consistent():
    d = json.load(open(os.path.join(ROOT, "profiles", "r1_v4_scene_build.json")))
    for key in ("c4_upload_dense", "c5_build_world", "c5_upload_dense"):
        e = d[key]
        assert abs(e["gbs"] - e["bytes"] / e["ms"] / 1e6) < 0.2
        assert abs(e["frac_of_hbm_peak"] - e["gbs"] / d["peak_gbs"]) < 1e-3
        assert 0.0 < e["frac_of_hbm_peak"] < 1.0
