#!/usr/bin/env python
"""bench.py — headline benchmark of the sphere-tracing render pass (BASELINE.json metric:
"Mrays/s and FPS at 4K on fixed CSG scene, 1/2/4/8 B200; % FP32 roofline").

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--mode fast|exact]
                  [--workload cfg3|cfg2|cfg1|cfg4|cfg5|n4] [--kernel wavefront|simple]

A "step" is one pass of the hot path over one batch: rm_paint of the workload's frame(s).
  cfg3 (default)  configs[2]: deep CSG scene S2 (64 graph nodes, smooth-min) at 3840x2160 — the config
                  the metric is quoted on.  N>1: every rank renders its own 4K frame of an orbit-pose
                  batch per step (frame-parallel, no data-path collective)      -> "scaling": "weak"
  cfg4            configs[3]: ONE 7680x4320 S2 frame per step, interleaved 16-row strips over the ranks,
                  packed bands gathered to rank 0 over NCCL and re-ordered        -> "scaling": "strong"
  cfg5            configs[4]: 1024 orbit poses of S1 at 1080p per step, poses sharded by frame -> "strong"
  cfg1 / cfg2     configs[0] / configs[1] single-GPU parity-sized cases
Prints ONE JSON line on rank 0.
"""
from __future__ import annotations

import argparse
import importlib
import json
import os
import subprocess
import sys
import threading
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

WORKLOADS = {
    "cfg1": ("S0", 640, 480, 1, "configs[0]: single-sphere S0 640x480"),
    "cfg2": ("S1", 1920, 1080, 1, "configs[1]: editor CSG scene S1 (box U sphere \\ sphere) 1920x1080"),
    "cfg3": ("S2", 3840, 2160, 1, "configs[2]: deep CSG S2 (32 leaves, 28 smooth-unions, 3 subtractions; 251 words) 3840x2160"),
    "cfg4": ("S2", 7680, 4320, 1, "configs[3]: S2 at 7680x4320, one frame row-sharded in interleaved 16-row strips, NCCL gather to rank 0"),
    "cfg5": ("S1", 1920, 1080, 1024, "configs[4]: 1024 orbit poses of S1 at 1920x1080, frame-parallel"),
    # not a BASELINE.json config: the N4 extension scene (plane + translation / rotation / scale groups), single GPU
    "n4": ("S3", 1920, 1080, 1, "extension (SURVEY 8(f) N4, not a BASELINE config): S3, 8 rotated / translated / scaled instances over a plane, 1920x1080"),
}
F_STEP, F_HIT, F_SAMPLE, F_FLOOR, F_PIXEL = 9, 75, 87, 14, 31  # SURVEY.md §8(a) FLOP table
OP_FLOPS = {0: (10, 4), 1: (22, 6), 100: (1, 0), 101: (2, 0), 102: (1, 0), 103: (10, 1),  # opcode: (flops, params)
            # N4, counted the same literal way: plane 3 mul + 3 add, translation 3 sub, rotation 9 mul + 6 add, scale 3 div + 1 mul
            2: (6, 4), 200: (3, 3), 201: (0, 0), 202: (15, 9), 203: (0, 0), 204: (3, 1), 205: (1, 1)}


def f_map(words, cmd_count) -> int:
    i = fl = 0
    for _ in range(cmd_count):
        f, n = OP_FLOPS[int(words[i])]
        fl += f
        i += 1 + n
    return fl


def frame_flops(counters, fmap, n_pixels) -> float:
    n_step, n_hit, n_floor, n_sample = (float(x) for x in counters)
    return (n_step + 4 * n_hit) * fmap + F_STEP * n_step + F_HIT * n_hit + F_SAMPLE * n_sample + F_FLOOR * n_floor + F_PIXEL * n_pixels


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
            self._stop.wait(0.05)

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


def ncu_traffic(workload: str, kernel: str, specialised: bool = False):
    """DRAM bytes per launch of the dominant kernel from the committed `ncu --set full` capture of this exact
    workload + kernel (profiles/*.ncu_summary.csv), or None.  Not measured live: a profiler is never in a timed run."""
    name = {("cfg3", "wavefront2", False): "r1_s3_wavefront2_R8_cfg3.ncu_summary.csv",
            ("cfg3", "direct", False): "r1_final_direct_R8_cfg3.ncu_summary.csv",
            ("cfg3", "direct", True): "r1_s14_spec_R2_cfg3.ncu_summary.csv",
            ("n4", "direct", True): "r1_s17_spec_R4_s3.ncu_summary.csv"}.get((workload, kernel, specialised))
    if not name or not (ROOT / "profiles" / name).exists():
        return None, None
    tot, unit_scale = 0.0, {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    for line in (ROOT / "profiles" / name).read_text().splitlines():
        parts = [p.strip('"') for p in line.split('","')]
        if len(parts) == 3 and parts[0] in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
            tot += float(parts[1].replace(",", "")) * unit_scale.get(parts[2], 1.0)
    return (tot or None), f"profiles/{name}"


def cpu_sample_rows(H: int, n: int):
    n = max(1, min(n, H))
    return sorted({min(H - 1, int((i + 0.5) * H / n)) for i in range(n)})  # evenly spaced over the whole frame


def run_cpu_oracle(cmd, words, W, H, uniforms, rows):
    """Times the CPU oracle (all host threads) on single rows of the workload frame.  The oracle is
    test infrastructure: this is the cpu_baseline / --impl reference leg, never the product path."""
    sys.path.insert(0, str(ROOT / "tests"))
    import helpers as h
    nthreads = h.oracle_lib().rmo_num_threads()
    t0 = time.perf_counter()
    for r in rows:
        h.oracle_render(uniforms, cmd, words, W, H, row0=r, nrows=1)
    dt = time.perf_counter() - t0
    return len(rows) * W * 16 / dt / 1e6, nthreads, dt


def measure_cpu_baseline(cmd, words, W, H, uniforms, scene, first_rows: int = 108, target_s: float = 14.0) -> dict:
    """The cpu_baseline object of the JSON line: the CPU oracle port on all host threads, on evenly spaced rows of the
    workload frame.  A first pass of `first_rows` rows sizes the sample for 10-30 s of CPU work on this box."""
    rows = cpu_sample_rows(H, first_rows)
    v, nthreads, dt = run_cpu_oracle(cmd, words, W, H, uniforms, rows)
    if dt < 10.0 and len(rows) < H:
        rows = cpu_sample_rows(H, min(H, max(len(rows) + 1, int(len(rows) * target_s / max(dt, 0.05)))))
        v, nthreads, dt = run_cpu_oracle(cmd, words, W, H, uniforms, rows)
    return {"value": v, "unit": "Mrays/s", "cores": nthreads, "kind": "port",
            "sample": f"{len(rows)} evenly spaced rows of the {W}x{H} {scene} frame ({len(rows) * W * 16} rays, {dt:.1f} s)"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--mode", default="fast", choices=["fast", "exact"])
    ap.add_argument("--kernel", default="direct", choices=["wavefront", "simple", "wavefront2", "direct"])
    ap.add_argument("--workload", default="cfg3", choices=list(WORKLOADS))
    ap.add_argument("--specialize", default="on", choices=["on", "off"],
                    help="on (default): the scene's command stream is compiled into the kernel once (NVRTC, outside the timed "
                         "region, like the reference's shader compile at start-up); off: the decoded-program interpreter")
    ap.add_argument("--early-out", action="store_true",
                    help="opt-in exact early-out of the interpreter (bit-identical output, but NOT the shader's literal arithmetic "
                         "per evaluation: reported separately, never the headline)")
    ap.add_argument("--e2e", default="serial", choices=["serial", "pipelined"],
                    help="end-to-end leg of the single-frame workloads.  serial (default): rm_prepare + rm_paint + rm_download per step, "
                         "each step waits for its frame.  pipelined: the steps go through rm_paint_poses_host, which copies frame k to "
                         "host memory while frame k+1 renders (every step still uploads its uniforms and returns its frame)")
    ap.add_argument("--gather", default="p2p", choices=["p2p", "nccl"],
                    help="cfg4, N>1: p2p = every rank stores its strips straight into rank 0's frame over NVLink "
                         "(rm_paint_strips_into); nccl = packed bands + NCCL gather + re-order on rank 0")
    ap.add_argument("--poses", type=int, default=0, help="cfg5: override the 1024-pose batch size")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()

    import rm_import
    rm = rm_import.load()
    scenes = importlib.import_module("ray_marching_b200.scenes")
    sharding = importlib.import_module("ray_marching_b200.sharding")
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    scene, W, H, n_poses, desc = WORKLOADS[args.workload]
    if args.workload == "cfg5" and args.poses:
        n_poses = args.poses
    cmd, words = scenes.program(scene)  # flattened by the C++ host (BuildCommands)
    warm = max(args.warmup, 3)
    yaw0, pitch = (0.0, 0.0) if scene == "S0" else (0.6, 0.4)
    strong = args.workload in ("cfg4", "cfg5")
    config = {"workload": desc, "scene": scene, "width": W, "height": H, "samples_per_pixel": 16, "frames_per_step": n_poses if strong else world,
              "limits": [0.01, 100.0, 100], "l2": "flushed between timed steps (256 MiB memset, untimed)"}

    # ------------------------------------------------------------------ reference arm (CPU oracle)
    if args.impl == "reference":
        if rank != 0:
            return 0
        u = scenes.orbit_uniforms(W, H, yaw=yaw0, pitch=pitch)
        rows = cpu_sample_rows(H, 54)
        vals = []
        for s in range(warm + args.steps):
            v, nthreads, dt = run_cpu_oracle(cmd, words, W, H, u, rows)
            if s >= warm:
                vals.append((v, dt))
            if s == 0 and dt * (warm + args.steps) > 240:  # keep the whole run within a few minutes
                rows = rows[: max(4, int(len(rows) * 240 / (dt * (warm + args.steps))))]
        v = float(np.mean([x[0] for x in vals]))
        ms = float(np.mean([x[1] for x in vals])) * 1e3
        sample = f"{len(rows)} evenly spaced rows of the {W}x{H} {scene} frame per step ({len(rows) * W * 16} rays)"
        print(json.dumps({
            "impl": "reference", "metric": "Mrays/s", "value": v, "unit": "Mrays/s", "n_gpus": args.gpus, "steps": args.steps,
            "warmup": warm, "ms_per_step": ms, "higher_is_better": True, "scaling": "strong" if strong else "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": config,
            "cpu_baseline": {"value": v, "unit": "Mrays/s", "cores": nthreads, "kind": "port", "sample": sample},
            "e2e": {"value": v, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "note": "reference arm = CPU oracle port of ray_marching.wgsl; the Rust/WGSL reference itself cannot run on this box "
                    "(no rustc, no graphics stack: wgpu-on-lavapipe unavailable)"}))
        return 0

    # ------------------------------------------------------------------ B200 arm
    import torch
    dist = None
    torch.cuda.set_device(local_rank)
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    dev = torch.device("cuda", local_rank)
    ctx = rm.Context(local_rank)
    ctx.set_mode(rm.MODE_FAST if args.mode == "fast" else rm.MODE_EXACT)
    ctx.set_kernel({"wavefront": rm.KERNEL_WAVEFRONT, "simple": rm.KERNEL_SIMPLE, "wavefront2": rm.KERNEL_WAVEFRONT2, "direct": rm.KERNEL_DIRECT}[args.kernel])
    # a dedicated non-default stream: the ABI reads a NULL stream as "the ctx's own stream", and the
    # CUDA events that time the steps must sit on the stream the kernels are launched on
    if args.early_out:
        ctx.set_early_out(True)
        config["early_out"] = "ON: executed work per map_scene evaluation is less than the algorithmic FLOPs the roofline fraction counts"
    stream = torch.cuda.Stream(dev)
    torch.cuda.set_stream(stream)
    sp = stream.cuda_stream
    assert sp != 0

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize(dev)
        ctx.sync()

    # ---- per-workload step --------------------------------------------------------------------
    if args.workload == "cfg4":
        strip = sharding.DEFAULT_STRIP_ROWS
        rows_max = sharding.max_local_rows(H, strip, world)
        band = torch.zeros((rows_max, W, 4), dtype=torch.uint8, device=dev)
        gathered = [torch.zeros_like(band) for _ in range(world)] if (rank == 0 and world > 1) else None
        uniforms = scenes.orbit_uniforms(W, H, yaw=yaw0, pitch=pitch)
        ctx.prepare(uniforms, cmd, words)
        rays_local = sharding.local_rows(H, strip, rank, world) * W * 16
        rays_step_total = W * H * 16
        frame_out = {}
        p2p = dist is not None and args.gather == "p2p"
        if p2p:
            # rank 0 owns the frame; the others map it (CUDA IPC) and their render kernels store into it over NVLink
            hbuf = torch.zeros(64, dtype=torch.uint8, device=dev)
            if rank == 0:
                frame_ptr, hnd = ctx.shared_frame_create(W * H * 4)
                hbuf.copy_(torch.frombuffer(bytearray(hnd), dtype=torch.uint8))
            dist.broadcast(hbuf, src=0)
            if rank != 0:
                frame_ptr = ctx.shared_frame_open(bytes(hbuf.cpu().numpy().tobytes()))
            done = torch.zeros(1, device=dev)
            config.update(sharding=f"interleaved strips of {strip} rows, strip s -> rank s % N; every rank's render kernel stores its "
                                   "strips straight into rank 0's frame over NVLink (CUDA IPC mapping), completion = one tiny all-reduce")

            def step():
                ctx.paint_strips_into(W, H, strip, rank, world, frame_ptr, W * 4, sp)
                dist.all_reduce(done)  # enqueued behind the paint on every rank: rank 0 passes it once all strips are stored
                frame_out["ptr"] = frame_ptr
        else:
            config.update(sharding=f"interleaved strips of {strip} rows, strip s -> rank s % N; NCCL gather + indexed re-order on rank 0")

            def step():
                ctx.paint_strips(W, H, strip, rank, world, band.data_ptr(), W * 4, sp)
                if dist is not None:
                    dist.gather(band, gathered, dst=0)
                    if rank == 0:
                        frame_out["f"] = sharding.reassemble(torch.stack(gathered), H, strip, world)
                else:
                    frame_out["f"] = band

        def count_step():
            ctx.paint_strips(W, H, strip, rank, world, band.data_ptr(), W * 4, sp)
        n_pixels_local = rays_local // 16
    elif args.workload == "cfg5":
        my = list(sharding.pose_slice(n_poses, rank, world))
        all_u = scenes.orbit_pose_batch(W, H, n_poses, pitch=0.4)
        my_u = b"".join(all_u[k * 144:(k + 1) * 144] for k in my)
        ctx.prepare(all_u[:144], cmd, words)
        rays_local = len(my) * W * H * 16
        rays_step_total = n_poses * W * H * 16
        config.update(sharding="pose k -> rank k % N; frames stay on their GPU (throughput mode)", poses=n_poses)

        def step():
            ctx.paint_poses(my_u, len(my), W, H, None, 0, 0, sp)
        count_step = step
        n_pixels_local = len(my) * W * H
    else:
        # weak scaling = FIXED work per GPU: every rank renders the same configs[2] pose.  (A different yaw per rank changes
        # the ray-march work per rank; measured on 8 GPUs the costliest of 8 yaws took 47.19 ms against 44.01 ms for this
        # pose, which is view imbalance, not scaling loss — profiles/r1_s13_NOTES.md.)
        uniforms = scenes.orbit_uniforms(W, H, yaw=yaw0, pitch=pitch)
        ctx.prepare(uniforms, cmd, words)
        rays_local = W * H * 16
        rays_step_total = rays_local * world
        config.update(sharding="one frame per rank per step (the same pose on every rank: fixed work per GPU), no collective")

        def step():
            ctx.paint(W, H, 0, H, None, 0, sp)
        count_step = step
        n_pixels_local = W * H

    # program specialisation: compile map_scene of this command stream into the kernel (once, untimed; cached process-wide)
    spec = {"requested": args.specialize == "on" and args.kernel == "direct" and not args.early_out, "active": False}
    if spec["requested"]:
        try:
            ctx.specialize()
        except rm.RmError as e:  # no NVRTC on the box, or the program is too long: the interpreter kernel runs, and the line says so
            spec["reason"] = str(e)[:300]
        spec.update({k: v for k, v in ctx.spec_info().items() if k in ("active", "rays_per_thread", "min_blocks", "compile_ms", "cubin_bytes")})
        spec["note"] = ("map_scene of this command stream compiled into the kernel at run time (NVRTC), once per program and outside the timed "
                        "region, like the reference's shader compile at start-up; same arithmetic per ray, exact mode bit-identical to the oracle")

    # work counters of this rank's step (untimed) -> algorithmic FLOPs
    ctx.set_counters(True)
    count_step()
    counters = ctx.get_counters()
    ctx.set_counters(False)
    fmap = f_map(words, cmd)
    flops = frame_flops(counters, fmap, n_pixels_local)
    peak = ctx.fp32_peak()

    for _ in range(warm):
        step()
    barrier()
    launches0 = ctx.launch_count
    spec_launches0 = ctx.spec_info()["launches"]
    ev0 = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps)]
    ev1 = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps)]
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    with ClockSampler(local_rank) as clk:
        t_wall0 = time.perf_counter()
        for i in range(args.steps):
            flush.zero_()                 # L2 flush, outside the timed event pair
            ev0[i].record(stream)
            step()
            ev1[i].record(stream)
        barrier()
        t_wall = time.perf_counter() - t_wall0
    launches = ctx.launch_count - launches0
    per_step = [a.elapsed_time(b) for a, b in zip(ev0, ev1)]
    ms_local = float(np.mean(per_step))
    if spec["active"]:
        spec["launches_in_timed_region"] = ctx.spec_info()["launches"] - spec_launches0
        # the same steps through the interpreter kernel, for the record (same harness, a few steps)
        ctx.set_specialize(False)
        step()
        barrier()
        n_i = max(2, min(args.steps, 4))
        iv0 = [torch.cuda.Event(enable_timing=True) for _ in range(n_i)]
        iv1 = [torch.cuda.Event(enable_timing=True) for _ in range(n_i)]
        for i in range(n_i):
            flush.zero_()
            iv0[i].record(stream)
            step()
            iv1[i].record(stream)
        barrier()
        spec["interpreter_ms_per_step_this_rank"] = float(np.mean([a.elapsed_time(b) for a, b in zip(iv0, iv1)]))
        ctx.specialize()  # back on (a cache hit) for the end-to-end leg

    # ---- e2e: host buffers in, host result out, through the public C-ABI calls (cfg1-3: prepare + paint + download)
    e2e_ms_local, h2d, d2h = None, 0, 0
    if not strong and args.e2e == "pipelined":
        n_e2e = max(3, min(args.steps, 10))
        ring_frames = 3
        pinned = rm.PinnedBuffer(ring_frames * W * H * 4)
        batch = uniforms * n_e2e                   # one 144-byte block per step, in host memory
        ctx.prepare(uniforms, cmd, words)
        ctx.paint_poses_host(batch[:2 * 144], 2, W, H, pinned.ptr, W * H * 4, ring_frames)  # warm-up (allocates the device ring)
        barrier()
        t0 = time.perf_counter()
        up0 = ctx.program_uploads
        ctx.prepare(uniforms, cmd, words)          # unchanged words: only the uniforms are replaced
        ctx.paint_poses_host(batch, n_e2e, W, H, pinned.ptr, W * H * 4, ring_frames)  # returns with every frame in host memory
        e2e_ms_local = (time.perf_counter() - t0) * 1e3 / n_e2e
        h2d, d2h = 144 + 12 + (ctx.program_uploads - up0) * (4 + int(len(words)) * 4) // n_e2e, W * H * 4
        config["e2e_mode"] = "pipelined: rm_paint_poses_host, frame k copied to host while frame k+1 renders"
        pinned.free()
    elif not strong:
        pinned = rm.PinnedBuffer(W * H * 4)
        n_e2e = max(3, min(args.steps, 10))
        for it in range(2 + n_e2e):
            if it == 2:
                barrier()
                t0 = time.perf_counter()
                up0 = ctx.program_uploads
            ctx.prepare(uniforms, cmd, words)      # H2D: uniforms + decoded program
            ctx.paint(W, H)
            ctx.download(pinned)                   # D2H: the RGBA8 frame (synchronises)
        e2e_ms_local = (time.perf_counter() - t0) * 1e3 / n_e2e
        # uniforms + limits travel with every launch; the program only when rm_prepare really uploaded it
        h2d, d2h = 144 + 12 + (ctx.program_uploads - up0) * (4 + int(len(words)) * 4) // n_e2e, W * H * 4
        pinned.free()
    elif args.workload == "cfg4":
        host_frame = torch.empty((H, W, 4), dtype=torch.uint8).pin_memory() if rank == 0 else None
        n_e2e = max(3, min(args.steps, 5))
        for it in range(1 + n_e2e):
            if it == 1:
                barrier()
                t0 = time.perf_counter()
                up0 = ctx.program_uploads
            ctx.prepare(uniforms, cmd, words)
            step()
            if rank == 0:
                if "ptr" in frame_out:
                    stream.synchronize()
                    ctx.copy_to_host(frame_out["ptr"], host_frame.numpy(), W * H * 4, sp)
                else:
                    host_frame.copy_(frame_out["f"][:H], non_blocking=False)
            barrier()
        e2e_ms_local = (time.perf_counter() - t0) * 1e3 / n_e2e
        # uniforms + limits travel with every launch; the program only when rm_prepare really uploaded it
        h2d, d2h = 144 + 12 + (ctx.program_uploads - up0) * (4 + int(len(words)) * 4) // n_e2e, W * H * 4
        if rank == 0:  # untimed: the frame the ranks assembled must equal rank 0's own full-frame render, byte for byte
            alone = np.empty((H, W, 4), np.uint8)
            ctx.paint(W, H)
            ctx.download(alone)
            config["frame_check"] = bool(np.array_equal(alone, host_frame.numpy()))
        barrier()
    elif args.workload == "cfg5":
        # host uniform blocks in, EVERY frame back in host memory (rm_paint_poses_host: the D2H copy of frame p
        # overlaps the render of frame p+1); the frames land in a pinned 32-frame host ring, as a consumer would drain them
        ring_frames = 32
        pinned = rm.PinnedBuffer(ring_frames * W * H * 4)
        n_e2e = 2
        for it in range(1 + n_e2e):
            if it == 1:
                barrier()
                t0 = time.perf_counter()
                up0 = ctx.program_uploads
            ctx.prepare(all_u[:144], cmd, words)
            ctx.paint_poses_host(my_u, len(my), W, H, pinned.ptr, W * H * 4, ring_frames)
            barrier()
        e2e_ms_local = (time.perf_counter() - t0) * 1e3 / n_e2e
        h2d, d2h = n_poses * 144 + 12 + (ctx.program_uploads - up0) * (4 + int(len(words)) * 4) // n_e2e, n_poses * W * H * 4
        pinned.free()

    if dist is not None:
        tt = torch.tensor([ms_local, e2e_ms_local or 0.0], device=dev, dtype=torch.float64)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        ms, e2e_ms = float(tt[0]), float(tt[1])
    else:
        ms, e2e_ms = ms_local, e2e_ms_local or 0.0
    if rank == 0:
        value = rays_step_total / (ms * 1e-3) / 1e6
        achieved = flops / (ms_local * 1e-3) / 1e12  # rank 0's kernel work over rank 0's device time
        fb_bytes = n_pixels_local * 4
        traffic, traffic_src = ncu_traffic(args.workload, args.kernel, spec["active"]) if args.mode == "fast" and world == 1 else (None, None)
        roof = {"bound": "fp32", "achieved": achieved, "peak": peak["tflops"], "unit": "TFLOP/s", "frac": achieved / peak["tflops"],
                "traffic": traffic, "traffic_unit": "bytes of DRAM read+write per launch", "traffic_source": traffic_src,
                "peak_source": "FFMA micro-benchmark run live on this GPU (rm_fp32_peak); MEASURED_PEAKS.json has no FP32 ALU peak; "
                               "nominal 148 SM x 128 lanes x 2 x 1.965 GHz = 74.4",
                "algorithmic_flops_per_step": flops, "f_map_flops_per_eval": fmap,
                "counters": {"n_step": int(counters[0]), "n_hit": int(counters[1]), "n_floor": int(counters[2]), "n_sample": int(counters[3])},
                "framebuffer_gbs": fb_bytes / (ms_local * 1e-3) / 1e9,
                "hbm_frac_of_measured_6531.9": fb_bytes / (ms_local * 1e-3) / 1e9 / 6531.9}
        line = {"metric": "Mrays/s", "value": value, "unit": "Mrays/s", "n_gpus": world, "steps": args.steps, "warmup": warm,
                "ms_per_step": ms, "fps": (n_poses if strong else world) * 1e3 / ms, "higher_is_better": True,
                "scaling": "strong" if strong else "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "config": dict(config, mode=args.mode, kernel=args.kernel, specialize=spec), "clocks": clk.summary(), "gpu_launches": int(launches),
                "roofline": roof, "wall_s_timed_region": t_wall}
        if e2e_ms:
            line["e2e"] = {"value": rays_step_total / (e2e_ms * 1e-3) / 1e6, "unit": "Mrays/s", "ms_per_step": e2e_ms,
                           "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h}
        if not args.no_cpu_baseline and world == 1:
            line["cpu_baseline"] = measure_cpu_baseline(cmd, words, W, H, scenes.orbit_uniforms(W, H, yaw=yaw0, pitch=pitch), scene)
        print(json.dumps(line))
    ctx.close()
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
