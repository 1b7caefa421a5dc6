#!/usr/bin/env python
"""Times the N4 extension scene S3 (plane + translation / rotation / scale groups) on one GPU through the C ABI:
the EXT interpreter build of the production kernel against the run-time specialised build, fast mode, CUDA events on
the context's stream (rm_mark / rm_elapsed_ms), L2 flushed between frames.  No torch: starts in about a second.
Prints one JSON line per variant.  Not a BASELINE.json config — bench.py --workload n4 is the full-contract form.

  python tools/n4_timing.py [--width 1920 --height 1080 --frames 10] [--scene S3|short] [--lib other/librm_b200.so]
--scene short: Translation(SmoothUnion(Sphere, Plane)), 3 decoded instructions -> the 4-rays-per-thread variant.
--lib: time another build of the library (e.g. one compiled with -DRM_DIRECT_MINB8_EXT=...).
"""
import argparse
import importlib
import json
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import bench  # noqa: E402  (f_map / frame_flops: the FLOP model of SURVEY.md 8(a) + the N4 rows)
import rm_import  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--width", type=int, default=1920)
    ap.add_argument("--height", type=int, default=1080)
    ap.add_argument("--frames", type=int, default=10)
    ap.add_argument("--scene", default="S3")
    ap.add_argument("--lib", default=None)
    ap.add_argument("--variants", default="interpreter,specialised")
    a = ap.parse_args()
    rm = rm_import.load()
    if a.lib:
        rm.LIB_PATH = Path(a.lib).resolve()
    scenes = importlib.import_module("ray_marching_b200.scenes")
    host = importlib.import_module("ray_marching_b200.host")
    W, H = a.width, a.height
    if a.scene == "short":
        cmd, words = host.build_commands(host.translation(host.binary(host.SMOOTH_UNION, host.sphere((0, 0, 0), 1.0),
                                                                         host.plane((0, 1, 0), 0.8), 0.5), (0.5, 0.2, 0.0)))
    else:
        cmd, words = scenes.program(a.scene)
    u = scenes.orbit_uniforms(W, H, yaw=0.6, pitch=0.4)
    fmap = bench.f_map(words, cmd)
    with rm.Context(0) as ctx:
        peak = ctx.fp32_peak()
        ctx.set_mode(rm.MODE_FAST)
        ctx.prepare(u, cmd, words)
        ctx.set_counters(True)
        ctx.paint(W, H)
        ctx.sync()
        counters = ctx.get_counters()
        ctx.set_counters(False)
        flops = bench.frame_flops(counters, fmap, W * H)
        for variant in a.variants.split(","):
            if variant == "specialised":
                ctx.specialize()
            info = ctx.spec_info()
            for _ in range(3):
                ctx.paint(W, H)
            ms = []
            for _ in range(a.frames):
                ctx.flush_l2()
                ctx.mark(0)
                ctx.paint(W, H)
                ctx.mark(1)
                ctx.sync()
                ms.append(ctx.elapsed_ms(0, 1))
            m = float(np.median(ms))
            print(json.dumps({"scene": a.scene, "width": W, "height": H, "variant": variant, "ms_per_frame": m, "ms_min": min(ms),
                              "Mrays/s": W * H * 16 / (m * 1e-3) / 1e6, "f_map_flops_per_eval": fmap,
                              "frac_of_fp32_peak": flops / (m * 1e-3) / 1e12 / peak["tflops"], "fp32_peak_tflops": peak["tflops"],
                              "rays_per_thread": info["rays_per_thread"] if info["active"] else None,
                              "counters": [int(x) for x in counters], "cmd_count": int(cmd), "n_words": int(len(words)),
                              "lib": str(rm.LIB_PATH.name if not a.lib else a.lib)}))


if __name__ == "__main__":
    main()
