#!/usr/bin/env python
"""bench.py — headline benchmark of the sphere-tracing render pass (BASELINE.json metric).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--mode fast|exact]
                  [--workload cfg3|cfg2|cfg1] [--kernel wavefront|simple]

A "step" is one pass of the hot path over one frame: rm_paint of the workload's frame.  Workload at
N=1 = BASELINE.json configs[2]: the deep CSG scene S2 (64 graph nodes, smooth-min blends) at
3840x2160 on 1 B200 — the configuration the metric "Mrays/s and FPS at 4K" is quoted on.
N>1 (torchrun, one rank per GPU): every rank renders its own 4K frame of an orbit-pose batch
(frame-parallel, no data-path collective) -> "scaling": "weak".
Prints ONE JSON line on rank 0 (contract in the task statement).
"""
from __future__ import annotations

import argparse
import json
import math
import os
import subprocess
import sys
import threading
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))

WORKLOADS = {
    # name: (scene, W, H, description)
    "cfg1": ("S0", 640, 480, "configs[0]: single-sphere S0 640x480"),
    "cfg2": ("S1", 1920, 1080, "configs[1]: editor CSG scene S1 (box U sphere \\ sphere) 1920x1080"),
    "cfg3": ("S2", 3840, 2160, "configs[2]: deep CSG S2 (32 leaves, 28 smooth-unions, 3 subtractions; 251 words) 3840x2160"),
}
# algorithmic FLOP table, SURVEY.md §8(a)/(d)
F_STEP, F_HIT, F_SAMPLE, F_FLOOR, F_PIXEL = 9, 75, 87, 14, 31


def f_map(words, cmd_count) -> int:
    import helpers as h
    per = {h.OP_SPHERE: (10, 4), h.OP_BOX: (22, 6), h.OP_UNION: (1, 0), h.OP_SUBTRACTION: (2, 0),
           h.OP_INTERSECTION: (1, 0), h.OP_SMOOTH_UNION: (10, 1)}
    i = fl = 0
    for _ in range(cmd_count):
        f, n = per[int(words[i])]
        fl += f
        i += 1 + n
    return fl


def frame_flops(counters, fmap, W, H) -> float:
    n_step, n_hit, n_floor, n_sample = (float(x) for x in counters)
    n_eval = n_step + 4 * n_hit
    return n_eval * fmap + F_STEP * n_step + F_HIT * n_hit + F_SAMPLE * n_sample + F_FLOOR * n_floor + F_PIXEL * W * H


class ClockSampler:
    """Samples nvidia-smi SM clocks and throttle reasons during the timed region."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.rows, self._stop, self._t = index, [], threading.Event(), None

    def _run(self):
        while not self._stop.is_set():
            try:
                out = subprocess.run(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-i",
                                      str(self.index)], capture_output=True, text=True, timeout=5).stdout.strip()
                if out:
                    self.rows.append([x.strip() for x in out.split(",")])
            except Exception:
                pass
            self._stop.wait(0.1)

    def __enter__(self):
        self._t = threading.Thread(target=self._run, daemon=True)
        self._t.start()
        return self

    def __exit__(self, *a):
        self._stop.set()
        self._t.join(timeout=6)

    def summary(self):
        if not self.rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        sm = sorted(float(r[0]) for r in self.rows)
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(r[3 + i].lower().startswith("active") for r in self.rows)]
        pw = [float(r[2]) for r in self.rows if r[2].replace(".", "", 1).isdigit()]
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": float(self.rows[0][1]), "reasons": reasons,
                "samples": len(self.rows), "power_w_max": max(pw) if pw else None}


def cpu_sample_rows(H: int, n: int = 108):
    step = max(1, H // n)
    return list(range(step // 2, H, step))[:n]


def run_cpu_oracle(scene, W, H, uniforms, rows):
    """Times the CPU oracle (all host threads) on single rows of the workload frame."""
    import helpers as h
    cmd, words = h.SCENES[scene]()
    nthreads = h.oracle_lib().rmo_num_threads()
    t0 = time.perf_counter()
    for r in rows:
        h.oracle_render(uniforms, cmd, words, W, H, row0=r, nrows=1)
    dt = time.perf_counter() - t0
    rays = len(rows) * W * 16
    return rays / dt / 1e6, nthreads, dt


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--mode", default="fast", choices=["fast", "exact"])
    ap.add_argument("--kernel", default="wavefront", choices=["wavefront", "simple"])
    ap.add_argument("--workload", default="cfg3", choices=list(WORKLOADS))
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()

    import helpers as h
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    scene, W, H, desc = WORKLOADS[args.workload]
    cmd, words = h.SCENES[scene]()
    warm = max(args.warmup, 3)
    yaw0, pitch = (0.0, 0.0) if scene == "S0" else (0.6, 0.4)
    yaw = yaw0 + 2 * math.pi * rank / max(world, 1)
    uniforms = h.orbit_uniforms(W, H, yaw=yaw, pitch=pitch)
    rays_per_frame = W * H * 16
    config = {"workload": desc, "scene": scene, "width": W, "height": H, "samples_per_pixel": 16,
              "rays_per_step": rays_per_frame, "limits": [0.01, 100.0, 100], "pose": "orbit yaw=0.6+2pi*rank/N pitch=0.4 r=5",
              "sharding": "one frame per rank per step (frame-parallel), no collective", "l2": "flushed between timed steps (256 MiB memset, untimed)"}

    # ------------------------------------------------------------------ reference arm (CPU)
    if args.impl == "reference":
        if rank != 0:
            return 0
        rows = cpu_sample_rows(H, 54)
        vals = []
        for s in range(warm + args.steps):
            v, nthreads, dt = run_cpu_oracle(scene, W, H, h.orbit_uniforms(W, H, yaw=yaw0, pitch=pitch), rows)
            if s >= warm:
                vals.append((v, dt))
            if s == 0 and dt * (warm + args.steps) > 240:  # keep the whole run within a few minutes
                rows = rows[: max(4, int(len(rows) * 240 / (dt * (warm + args.steps))))]
        v = float(np.mean([x[0] for x in vals]))
        ms = float(np.mean([x[1] for x in vals])) * 1e3
        sample = f"{len(rows)} evenly spaced rows of the {W}x{H} {scene} frame per step ({len(rows) * W * 16} rays)"
        line = {"impl": "reference", "metric": "Mrays/s", "value": v, "unit": "Mrays/s", "n_gpus": args.gpus,
                "steps": args.steps, "warmup": warm, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": config,
                "cpu_baseline": {"value": v, "unit": "Mrays/s", "cores": nthreads, "kind": "port", "sample": sample},
                "e2e": {"value": v, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                "note": "reference = CPU oracle port of ray_marching.wgsl (the Rust/WGSL reference cannot run here: no rustc, no graphics stack)"}
        print(json.dumps(line))
        return 0

    # ------------------------------------------------------------------ B200 arm
    import torch
    import rm_import
    rm = rm_import.load()
    dist = None
    if world > 1:
        import torch.distributed as dist
        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    ctx = rm.Context(local_rank)
    ctx.set_mode(rm.MODE_FAST if args.mode == "fast" else rm.MODE_EXACT)
    ctx.set_kernel(rm.KERNEL_WAVEFRONT if args.kernel == "wavefront" else rm.KERNEL_SIMPLE)

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize(local_rank)
        ctx.sync()

    # work counters of this rank's frame (untimed) -> algorithmic FLOPs of one step
    ctx.prepare(uniforms, cmd, words)
    ctx.set_counters(True)
    ctx.paint(W, H)
    counters = ctx.get_counters()
    ctx.set_counters(False)
    fmap = f_map(words, cmd)
    flops = frame_flops(counters, fmap, W, H)
    peak = ctx.fp32_peak()

    for _ in range(warm):
        ctx.paint(W, H)
    barrier()
    launches0 = ctx.launch_count
    per_step = []
    with ClockSampler(local_rank) as clk:
        t_wall0 = time.perf_counter()
        for _ in range(args.steps):
            ctx.flush_l2()
            ctx.mark(0)
            ctx.paint(W, H)
            ctx.mark(1)
            per_step.append(ctx.elapsed_ms(0, 1))
        barrier()
        t_wall = time.perf_counter() - t_wall0
    launches = ctx.launch_count - launches0
    ms_local = float(np.mean(per_step))

    # e2e: host buffers in, host framebuffer out, through the C ABI (prepare + paint + download)
    pinned = rm.PinnedBuffer(W * H * 4)
    for _ in range(2):
        ctx.prepare(uniforms, cmd, words)
        ctx.paint(W, H)
        ctx.download(pinned)
    barrier()
    n_e2e = max(3, min(args.steps, 10))
    t0 = time.perf_counter()
    for _ in range(n_e2e):
        ctx.prepare(uniforms, cmd, words)
        ctx.paint(W, H)
        ctx.download(pinned)
    e2e_ms_local = (time.perf_counter() - t0) * 1e3 / n_e2e

    if dist is not None:
        tt = torch.tensor([ms_local, e2e_ms_local], device=f"cuda:{local_rank}", dtype=torch.float64)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        ms, e2e_ms = float(tt[0]), float(tt[1])
        ff = torch.tensor([flops], device=f"cuda:{local_rank}", dtype=torch.float64)
        dist.all_reduce(ff, op=dist.ReduceOp.SUM)
        flops_total = float(ff[0])
    else:
        ms, e2e_ms, flops_total = ms_local, e2e_ms_local, flops
    if rank == 0:
        total_rays = rays_per_frame * world
        value = total_rays / (ms * 1e-3) / 1e6
        e2e = total_rays / (e2e_ms * 1e-3) / 1e6
        achieved = flops / (ms_local * 1e-3) / 1e12  # this rank's kernel
        n_insn = 0
        roof = {"bound": "fp32", "achieved": achieved, "peak": peak["tflops"], "unit": "TFLOP/s",
                "frac": achieved / peak["tflops"], "traffic": None,
                "peak_source": "FFMA micro-benchmark run live on this GPU (rm_fp32_peak); MEASURED_PEAKS.json has no FP32 ALU peak",
                "algorithmic_flops_per_step": flops, "f_map_flops_per_eval": fmap,
                "counters": {"n_step": int(counters[0]), "n_hit": int(counters[1]), "n_floor": int(counters[2]), "n_sample": int(counters[3])},
                "framebuffer_gbs": W * H * 4 / (ms_local * 1e-3) / 1e9,
                "hbm_frac_of_measured": (W * H * 4 / (ms_local * 1e-3) / 1e9) / 6531.9}
        line = {"metric": "Mrays/s", "value": value, "unit": "Mrays/s", "n_gpus": world, "steps": args.steps, "warmup": warm,
                "ms_per_step": ms, "fps": 1e3 / ms * world, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": "f32", "data": "synthetic", "config": dict(config, mode=args.mode, kernel=args.kernel),
                "clocks": clk.summary(), "gpu_launches": int(launches),
                "e2e": {"value": e2e, "unit": "Mrays/s", "ms_per_step": e2e_ms, "h2d_bytes_per_step": 144 + 12 + 4 + int(len(words)) * 4,
                        "d2h_bytes_per_step": W * H * 4},
                "roofline": roof, "wall_s_timed_region": t_wall}
        if not args.no_cpu_baseline and world == 1:
            rows = cpu_sample_rows(H, 108)
            v, nthreads, dt = run_cpu_oracle(scene, W, H, uniforms, rows)
            line["cpu_baseline"] = {"value": v, "unit": "Mrays/s", "cores": nthreads, "kind": "port",
                                    "sample": f"{len(rows)} evenly spaced rows of the same {W}x{H} {scene} frame ({len(rows) * W * 16} rays, {dt:.1f} s)"}
        print(json.dumps(line))
    pinned.free()
    ctx.close()
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
