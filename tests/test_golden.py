"""Golden fixtures (tests/golden/*.npz, made by oracle/make_golden.py from the C oracle):
  * CPU: the C oracle still reproduces them, and the independent numpy restatement
    (oracle/np_oracle.py) agrees with them bit for bit;
  * GPU: RM_MODE_EXACT reproduces them bit for bit without needing the oracle at run time.
"""
import sys
from pathlib import Path

import numpy as np
import pytest

import helpers as h

ROOT = Path(__file__).resolve().parent.parent
GOLDEN = sorted((ROOT / "tests" / "golden").glob("*.npz"))


def load(p):
    z = np.load(p)
    return {k: z[k] for k in z.files}


@pytest.mark.parametrize("path", GOLDEN, ids=lambda p: p.stem)
def test_c_oracle_reproduces_golden(path):
    g = load(path)
    r = h.oracle_render(g["uniforms"].tobytes(), int(g["cmd_count"]), g["words"], int(g["W"]), int(g["H"]),
                        limits=g["limits"].tobytes(), aov=True)
    assert np.array_equal(r["rgba"], g["rgba"]) and np.array_equal(r["cls"], g["cls"])
    assert np.array_equal(r["t"].view(np.uint32), g["t"].view(np.uint32))
    assert np.array_equal(r["counters"], g["counters"])


@pytest.mark.parametrize("path", GOLDEN, ids=lambda p: p.stem)
def test_numpy_restatement_agrees_bit_for_bit(path):
    sys.path.insert(0, str(ROOT / "oracle"))
    import np_oracle
    g = load(path)
    lim = g["limits"].tobytes()
    limits = (*np.frombuffer(lim[:8], np.float32), int(np.frombuffer(lim[8:], np.uint32)[0]))
    r = np_oracle.render(g["uniforms"].tobytes(), int(g["cmd_count"]), g["words"], int(g["W"]), int(g["H"]), limits)
    assert np.array_equal(r["cls"], g["cls"])
    assert np.array_equal(r["rgba"], g["rgba"])
    hit = g["cls"] == 2
    assert np.array_equal(r["t"][hit], g["t"][hit])
    assert np.array_equal(r["counters"], g["counters"])


@pytest.mark.gpu
@pytest.mark.parametrize("path", GOLDEN, ids=lambda p: p.stem)
def test_gpu_exact_mode_reproduces_golden(path):
    import rm_import
    rm = rm_import.load()
    g = load(path)
    lim = g["limits"].tobytes()
    with rm.Context(0) as ctx:
        ctx.set_limits(*np.frombuffer(lim[:8], np.float32), int(np.frombuffer(lim[8:], np.uint32)[0]))
        ctx.set_mode(rm.MODE_EXACT)
        ctx.set_counters(True)
        img = ctx.render(g["uniforms"].tobytes(), int(g["cmd_count"]), g["words"], int(g["W"]), int(g["H"]))
        assert np.array_equal(img, g["rgba"])
        assert np.array_equal(ctx.get_counters(), g["counters"])
        dbg = ctx.paint_debug(int(g["W"]), int(g["H"]))
        assert np.array_equal(dbg["cls"], g["cls"])
        assert np.array_equal(dbg["t"].view(np.uint32), g["t"].view(np.uint32))
