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


def test_reference_arm_maps_nothing_of_the_product():
    """the CPU arm must not load librm_b200.so / librm_host.so — only the oracle."""
    code = ("import sys; sys.argv = ['bench.py', '--impl', 'reference', '--workload', 'cfg1', '--steps', '1', '--warmup', '3'];"
            f"sys.path.insert(0, {str(ROOT)!r}); import bench; bench.main();"
            "maps = open('/proc/self/maps').read();"
            "print('MAPPED', sorted({ln.split('/')[-1].strip() for ln in maps.splitlines() if 'librm_' in ln}));"
            "print('MODS', sorted(m for m in sys.modules if 'ray_marching' in m or m == 'rm_import' or m == 'torch'))")
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr[-2000:]
    assert "MAPPED ['librm_oracle.so']" in r.stdout, r.stdout[-500:]
    assert "MODS []" in r.stdout, r.stdout[-500:]
    d = json.loads([ln for ln in r.stdout.splitlines() if ln.startswith("{")][0])
    assert "ms_per_step_covers" in d and "bounded sample" in d["ms_per_step_covers"] and d["ms_whole_step_at_this_rate"] > 0


def test_parity_stats_gates():
    rng = np.random.default_rng(0)
    ref = rng.integers(0, 255, (3, 40, 4), dtype=np.uint8)
    cls = rng.integers(0, 3, (3, 40, 16), dtype=np.uint8)
    t = rng.uniform(1, 9, (3, 40, 16)).astype(np.float32)
    s = bench.parity_stats(ref.copy(), ref, cls.copy(), cls, t.copy(), t)
    assert s["rgba_identical"] and s["aov_identical"] and s["gates_ok"] and s["hitmask_pix_frac"] == 0 and s["t_rel_max"] == 0
    img = ref.copy()
    img[0, 0, 0] = (int(ref[0, 0, 0]) + 3) % 256 if ref[0, 0, 0] < 250 else ref[0, 0, 0] - 3
    cls2 = cls.copy()
    cls2[1, 5, 3] = 2 if cls[1, 5, 3] != 2 else 0
    t2 = t.copy()
    both = (cls == 2) & (cls2 == 2)
    idx = tuple(np.argwhere(both)[0])
    t2[idx] *= np.float32(1.01)
    s = bench.parity_stats(img, ref, cls2, cls, t2, t)
    assert s["rgba_max"] == 3 and abs(s["rgba_gt2_frac"] - 1 / 120) < 1e-9 and abs(s["hitmask_pix_frac"] - 1 / 120) < 1e-9
    assert 0.009 < s["t_rel_max"] < 0.011 and not s["gates_ok"] and not s["aov_identical"]


def test_hbm_peak_comes_from_measured_peaks_json():
    p = bench.measured_peaks()
    if (ROOT / "MEASURED_PEAKS.json").exists():
        assert p["hbm_gbs"] == json.loads((ROOT / "MEASURED_PEAKS.json").read_text())["hbm_gbs"]
    assert "6531.9" not in (ROOT / "bench.py").read_text()
