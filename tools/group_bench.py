#!/usr/bin/env python
"""Single-process multi-GPU timing of the rm_group_* surface (include/rm_b200.h): ONE host thread, N GPUs.
Supplements bench.py (whose N>1 arm is one process per GPU under torchrun, as the bench contract asks): same scenes,
same sizes, same kernel, but driven the way the reference's single render thread would drive them.

  python tools/group_bench.py --gpus 8 --workload cfg4        # one 8K frame, strips stored into GPU 0's frame
  python tools/group_bench.py --gpus 8 --workload cfg5        # 1024 orbit poses at 1080p, every frame returned to host

The calls are synchronous (they return when the frame is complete on GPU 0 / every frame is in host memory), so the
time of a step is the host's wall clock around the call: it includes the launches on every device and the final waits.
Prints one JSON line per run; not a bench.py replacement and not used by the driver."""
from __future__ import annotations

import argparse
import importlib
import json
import sys
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--workload", default="cfg4", choices=["cfg3", "cfg4", "cfg5"])
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--poses", type=int, default=1024)
    ap.add_argument("--chunk", type=int, default=256, help="cfg5: poses per rm_group_paint_poses call (pinned host slots)")
    ap.add_argument("--specialize", default="on", choices=["on", "off"], help="compile the scene into the kernel (rm_group_specialize)")
    ap.add_argument("--to-host", action="store_true", help="cfg3/cfg4: also copy the finished frame to pinned host memory")
    args = ap.parse_args()

    import rm_import
    rm = rm_import.load()
    scenes = importlib.import_module("ray_marching_b200.scenes")
    scene, W, H = {"cfg3": ("S2", 3840, 2160), "cfg4": ("S2", 7680, 4320), "cfg5": ("S1", 1920, 1080)}[args.workload]
    cmd, words = scenes.program(scene)
    g = rm.Group(list(range(args.gpus)))
    g.set_mode(rm.MODE_FAST)
    u = scenes.orbit_uniforms(W, H, yaw=0.6, pitch=0.4)
    g.prepare(u, cmd, words)
    if args.specialize == "on":
        g.specialize()  # one compile, shared by the members (process-wide cache)
    launches0 = [g.member_launch_count(i) for i in range(args.gpus)]
    times = []
    if args.workload == "cfg5":
        chunk = min(args.chunk, args.poses)
        assert args.poses % chunk == 0
        all_u = scenes.orbit_pose_batch(W, H, args.poses, pitch=0.4)
        buf = rm.PinnedBuffer(chunk * W * H * 4)

        def step():
            for c in range(args.poses // chunk):
                g.paint_poses(all_u[c * chunk * 144:(c + 1) * chunk * 144], chunk, W, H, buf.ptr)
        rays = args.poses * W * H * 16
        d2h = args.poses * W * H * 4
    else:
        buf = rm.PinnedBuffer(W * H * 4) if args.to_host else None

        def step():
            g.paint_bands(W, H, 16, buf)
        rays = W * H * 16
        d2h = W * H * 4 if args.to_host else 0
    for s in range(args.warmup + args.steps):
        t0 = time.perf_counter()
        step()
        dt = time.perf_counter() - t0
        if s >= args.warmup:
            times.append(dt)
        if s == args.warmup:
            launches0 = [g.member_launch_count(i) for i in range(args.gpus)]
    launches = sum(g.member_launch_count(i) - launches0[i] for i in range(args.gpus))
    ms = float(np.mean(times)) * 1e3
    print(json.dumps({
        "api": "rm_group (one host thread)", "workload": args.workload, "scene": scene, "width": W, "height": H, "n_gpus": args.gpus, "specialize": args.specialize,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "ms_min": float(np.min(times)) * 1e3,
        "value": rays / (ms * 1e-3) / 1e6, "unit": "Mrays/s", "d2h_bytes_per_step": d2h,
        "gpu_launches_after_first_timed_step": launches, "timing": "host wall clock around the synchronous call"}))
    g.close()


if __name__ == "__main__":
    main()
