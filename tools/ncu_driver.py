#!/usr/bin/env python
"""The smallest program that launches the kernel to be profiled: prepares one workload's frame and paints it a few times
through the C ABI (no torch).  Run it plain first, then under ncu (B200_PROFILING.md):

  python tools/ncu_driver.py --workload cfg3 --tier literal|structure|interp [--frames 4] [--mode fast|exact]
"""
import argparse
import importlib
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import bench  # noqa: E402
import rm_import  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="cfg3")
    ap.add_argument("--tier", default="literal", choices=["literal", "structure", "interp"])
    ap.add_argument("--frames", type=int, default=4)
    ap.add_argument("--mode", default="fast")
    ap.add_argument("--cull", action="store_true", help="opt-in per-tile culling (interpreter tier)")
    a = ap.parse_args()
    rm = rm_import.load()
    scenes = importlib.import_module("ray_marching_b200.scenes")
    scene, W, H, _, _ = bench.WORKLOADS[a.workload]
    cmd, words = scenes.program(scene)
    yaw, pitch = (0.0, 0.0) if scene == "S0" else (0.6, 0.4)
    u = scenes.orbit_uniforms(W, H, yaw=yaw, pitch=pitch)
    with rm.Context(0) as ctx:
        ctx.set_mode(rm.MODE_FAST if a.mode == "fast" else rm.MODE_EXACT)
        ctx.prepare(u, cmd, words)
        if a.cull:
            ctx.set_cull(True)
        if a.tier != "interp":
            ctx.set_spec_literal_after(8 if a.tier == "literal" else 0)
            ctx.specialize()
        for f in range(a.frames):
            ctx.mark(0)
            ctx.paint(W, H)
            ctx.mark(1)
            ctx.sync()
            print(f"frame {f}: {ctx.elapsed_ms(0, 1):.3f} ms  tier={ctx.spec_tier} info={ctx.spec_info()['rays_per_thread']}")
    return 0


if __name__ == "__main__":
    sys.exit(main())
