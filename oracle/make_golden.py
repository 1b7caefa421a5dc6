#!/usr/bin/env python
"""Generates tests/golden/*.npz from the C oracle (oracle/rm_oracle.c).  The reference has no golden
vectors of its own and cannot run here, so these fixtures pin the ORACLE (regression)
and give the GPU tests a reference that does not need the oracle at run time.
Usage: python oracle/make_golden.py
"""
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "tests"))
import helpers as h  # noqa: E402

CASES = [  # name, scene, W, H, yaw, pitch, limits
    ("s0_80x60", "S0", 80, 60, 0.0, 0.0, (0.01, 100.0, 100)),
    ("s1_96x54", "S1", 96, 54, 0.6, 0.4, (0.01, 100.0, 100)),
    ("s2_96x54", "S2", 96, 54, 0.6, 0.4, (0.01, 100.0, 100)),
    ("s2_50x30_limits", "S2", 50, 30, 2.2, -0.3, (0.02, 40.0, 24)),
    # N4 extension scene (plane + space transforms): kept in its own directory, read by tests/test_oracle_n4.py and
    # tests/test_n4_gpu.py
    ("n4/s3_96x54", "S3", 96, 54, 0.6, 0.4, (0.01, 100.0, 100)),
    ("n4/s3_64x40_limits", "S3", 64, 40, 2.2, 0.7, (0.02, 40.0, 24)),
]

if __name__ == "__main__":
    out = ROOT / "tests" / "golden"
    out.mkdir(exist_ok=True)
    for name, scene, W, H, yaw, pitch, lim in CASES:
        cmd, words = h.SCENES[scene]()
        u = h.orbit_uniforms(W, H, yaw=yaw, pitch=pitch)
        r = h.oracle_render(u, cmd, words, W, H, limits=h.limits_bytes(*lim), aov=True)
        (out / f"{name}.npz").parent.mkdir(exist_ok=True)
        np.savez_compressed(out / f"{name}.npz", uniforms=np.frombuffer(u, np.uint8), cmd_count=np.uint32(cmd),
                            words=words, W=np.uint32(W), H=np.uint32(H), limits=np.frombuffer(h.limits_bytes(*lim), np.uint8),
                            rgba=r["rgba"], cls=r["cls"], t=r["t"], counters=r["counters"])
        print(name, r["counters"])
