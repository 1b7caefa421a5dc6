#!/usr/bin/env python
"""Times the interpreter build of the production kernel on one workload through the C ABI (no torch; CUDA events on the
context's stream, L2 flushed between frames).  RM_INTERP_VARIANT / RM_DIRECT_R select the measurement variants of
rm_direct.cu; --lib times another build of the library.  One JSON line.

  [RM_INTERP_VARIANT=8t|16|16t] python tools/interp_timing.py [--workload cfg3] [--frames 6] [--lib variants/librm_x3.so]
"""
import argparse
import importlib
import json
import os
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import bench  # noqa: E402
import rm_import  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="cfg3")
    ap.add_argument("--frames", type=int, default=6)
    ap.add_argument("--lib", default=None)
    ap.add_argument("--tier", default="interp", choices=["interp", "structure", "literal"])
    a = ap.parse_args()
    rm = rm_import.load()
    if a.lib:
        rm.LIB_PATH = Path(a.lib).resolve()
    scenes = importlib.import_module("ray_marching_b200.scenes")
    scene, W, H, _, _ = bench.WORKLOADS[a.workload]
    cmd, words = scenes.program(scene)
    yaw, pitch = (0.0, 0.0) if scene == "S0" else (0.6, 0.4)
    u = scenes.orbit_uniforms(W, H, yaw=yaw, pitch=pitch)
    fmap = bench.f_map(words, cmd)
    with rm.Context(0) as ctx:
        peak = ctx.fp32_peak()
        ctx.set_mode(rm.MODE_FAST)
        ctx.prepare(u, cmd, words)
        if a.tier != "interp":
            ctx.set_spec_literal_after(8 if a.tier == "literal" else 0)
            ctx.specialize()
        ctx.set_counters(True)
        ctx.paint(W, H)
        cnt = ctx.get_counters()
        ctx.set_counters(False)
        ctx.paint(W, H)
        ctx.sync()
        ms = []
        for _ in range(a.frames):
            ctx.flush_l2()
            ctx.mark(0)
            ctx.paint(W, H)
            ctx.mark(1)
            ctx.sync()
            ms.append(ctx.elapsed_ms(0, 1))
        img = np.empty((H, W, 4), np.uint8)
        ctx.download(img)
        flops = bench.frame_flops(cnt, fmap, W * H)
        m = float(np.mean(ms))
        out = {"workload": a.workload, "tier": a.tier, "spec_tier": ctx.spec_tier, "fused": os.environ.get("RM_SPEC_FUSED", ""),
               "minb": os.environ.get("RM_SPEC_MINB", ""), "noinline": os.environ.get("RM_SPEC_NOINLINE", ""), "spec_r": os.environ.get("RM_SPEC_R", ""), "variant": os.environ.get("RM_INTERP_VARIANT", ""), "direct_r": os.environ.get("RM_DIRECT_R", ""),
               "lib": str(rm.LIB_PATH.name), "ms": m, "ms_min": float(np.min(ms)), "frac": flops / (m * 1e-3) / 1e12 / peak["tflops"],
               "sum": int(img.astype(np.uint64).sum())}
        print(json.dumps(out))
    return 0


if __name__ == "__main__":
    sys.exit(main())
