"""bench.py pieces that run without a GPU: the FLOP model, the CPU-baseline sampling and the reference arm's JSON line."""
import json
import subprocess
import sys
from pathlib import Path

import numpy as np
import pytest

import helpers as h

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import bench  # noqa: E402


@pytest.mark.parametrize("H,n", [(2160, 108), (2160, 54), (2160, 560), (1080, 2000), (480, 7), (5, 3)])
def test_sample_rows_are_unique_in_range_and_spread_over_the_whole_frame(H, n):
    rows = bench.cpu_sample_rows(H, n)
    assert rows == sorted(set(rows)) and rows[0] >= 0 and rows[-1] < H and len(rows) == min(n, H)
    assert rows[0] < H / len(rows) and rows[-1] >= H - H / len(rows) - 1


def test_flop_model_matches_the_survey_table_and_the_product_decoder():
    c, w = h.scene_s2()
    assert bench.f_map(w, c) == 798          # 16*10 + 16*22 + 28*10 + 3*2
    c, w = h.scene_s1()
    assert bench.f_map(w, c) == 22 + 10 + 1 + 10 + 2
    c, w = h.scene_s3()
    assert bench.f_map(w, c) == 461          # N4 rows: plane 6, translation 3, rotation 15, scale 3 + 1
    # the product's own count (Program::flops_per_eval, through the host-compiled decoder of tests/hostcheck)
    import test_hostcheck_cpu as t
    lib = t.load_hostcheck()
    for scene in (h.scene_s0, h.scene_s1, h.scene_s2, h.scene_s3):
        c, w = scene()
        assert t.hc_decode(lib, c, w)[3] == bench.f_map(w, c)


def test_cpu_baseline_object():
    c, w = h.scene_s1()
    u = h.orbit_uniforms(160, 90, yaw=0.6, pitch=0.4)
    b = bench.measure_cpu_baseline(c, w, 160, 90, u, "S1", first_rows=6, target_s=0.3)
    assert set(b) == {"value", "unit", "cores", "kind", "sample"}
    assert b["kind"] == "port" and b["unit"] == "Mrays/s" and b["value"] > 0 and b["cores"] >= 1
    assert "rows of the 160x90 S1 frame" in b["sample"]


def test_reference_arm_prints_one_contract_line():
    r = subprocess.run([sys.executable, str(ROOT / "bench.py"), "--impl", "reference", "--workload", "cfg1", "--steps", "1", "--warmup", "3"],
                       capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [ln for ln in r.stdout.splitlines() if ln.startswith("{")]
    assert len(lines) == 1
    d = json.loads(lines[0])
    for k in ("impl", "metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling", "vs_baseline",
              "dtype", "data", "config", "cpu_baseline", "e2e"):
        assert k in d, k
    assert d["impl"] == "reference" and d["unit"] == "Mrays/s" and d["value"] > 0 and d["vs_baseline"] is None
    assert d["e2e"] == {"value": d["value"], "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert d["cpu_baseline"]["kind"] == "port" and d["cpu_baseline"]["value"] == d["value"] and "workload" in d["config"]
