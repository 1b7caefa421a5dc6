#!/usr/bin/env python
"""bench.py — headline benchmark of the sphere-tracing render pass (BASELINE.json metric:
"Mrays/s and FPS at 4K on fixed CSG scene, 1/2/4/8 B200; % FP32 roofline").

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--mode fast|exact]
                  [--workload cfg3|cfg2|cfg1|cfg4|cfg5|n4] [--kernel direct|wavefront2|wavefront|simple]

A "step" is one pass of the hot path over one batch: rm_paint of the workload's frame(s).
  cfg3 (default)  configs[2]: deep CSG scene S2 (64 graph nodes, smooth-min) at 3840x2160 — the config the metric is
                  quoted on.  N>1: every rank renders the same 4K pose per step (fixed work per GPU, no data-path
                  collective) -> "scaling": "weak".  The same invocation then measures the two SHARDED configs of the
                  north star and reports them under "strong" (--strong off skips them):
                    cfg4  ONE 7680x4320 S2 frame per step, interleaved strips (16 rows, or the largest height that gives every rank the same count) over the ranks, every rank's render
                          kernel storing its strips straight into rank 0's frame over NVLink (CUDA IPC mapping)
                    cfg5  1024 orbit poses of S1 at 1080p per step, pose k -> rank k % N
                  each with the time of rank 0 doing the whole step ALONE on the same box (n1_ms_same_box), so that the
                  1 -> N ratio comes from one run on one box.
  cfg4 / cfg5     the same sharded configs as the headline of their own line              -> "scaling": "strong"
  cfg1 / cfg2     configs[0] / configs[1] single-GPU parity-sized cases;  n4: the extension scene S3
Prints ONE JSON line on rank 0.

--impl reference times the CPU oracle (oracle/, the reference's shader restated in C: the Rust/WGSL reference itself
cannot run on this box) and loads nothing of the product: scenes and uniforms come from tests/helpers.py.
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
    "cfg4": ("S2", 7680, 4320, 1, "configs[3]: S2 at 7680x4320, one frame row-sharded in interleaved strips, gathered on rank 0"),
    "cfg5": ("S1", 1920, 1080, 1024, "configs[4]: 1024 orbit poses of S1 at 1920x1080, frame-parallel"),
    # not a BASELINE.json config: the N4 extension scene (plane + translation / rotation / scale groups), single GPU
    "n4": ("S3", 1920, 1080, 1, "extension (SURVEY 8(f) N4, not a BASELINE config): S3, 8 rotated / translated / scaled instances over a plane, 1920x1080"),
}
F_STEP, F_HIT, F_SAMPLE, F_FLOOR, F_PIXEL = 9, 75, 87, 14, 31  # SURVEY.md §8(a) FLOP table
OP_FLOPS = {0: (10, 4), 1: (22, 6), 100: (1, 0), 101: (2, 0), 102: (1, 0), 103: (10, 1),  # opcode: (flops, params)
            # N4, counted the same literal way: plane 3 mul + 3 add, translation 3 sub, rotation 9 mul + 6 add, scale 3 div + 1 mul
            2: (6, 4), 200: (3, 3), 201: (0, 0), 202: (15, 9), 203: (0, 0), 204: (3, 1), 205: (1, 1)}
# north-star parity gates (BASELINE.json)
GATE_HITMASK, GATE_RGBA_LSB, GATE_RGBA_FRAC, GATE_T_REL, GATE_T_FRAC = 1e-4, 2, 1e-3, 1e-4, 1e-3


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


def measured_peaks() -> dict:
    """MEASURED_PEAKS.json (driver-written) or {}."""
    try:
        return json.loads((ROOT / "MEASURED_PEAKS.json").read_text())
    except Exception:
        return {}


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


# committed `ncu --set full` captures of (workload, kernel, specialised), newest first: DRAM bytes per launch
NCU_SUMMARIES = {("cfg3", "wavefront2", False): ["r1_s3_wavefront2_R8_cfg3.ncu_summary.csv"],
                 ("cfg3", "direct", False): ["r2_s47_interp_cfg3.ncu_summary.csv", "r2_s41_interp_cfg3.ncu_summary.csv", "r2_s30_interp_cfg3.ncu_summary.csv", "r2_s22_interp_cfg3.ncu_summary.csv", "r2_interp_cfg3.ncu_summary.csv", "r1_final_direct_R8_cfg3.ncu_summary.csv"],
                 ("cfg3", "direct", True): ["r2_s47_literal_cfg3.ncu_summary.csv", "r2_s41_literal_cfg3.ncu_summary.csv", "r2_s30_literal_cfg3.ncu_summary.csv", "r2_s22_literal_cfg3.ncu_summary.csv", "r2_spec_cfg3.ncu_summary.csv", "r1_s14_spec_R2_cfg3.ncu_summary.csv"],
                 ("n4", "direct", True): ["r1_s17_spec_R4_s3.ncu_summary.csv"]}


def ncu_traffic(workload: str, kernel: str, specialised: bool = False):
    """DRAM bytes per launch of the dominant kernel from the newest committed `ncu --set full` capture of this exact
    workload + kernel (profiles/*.ncu_summary.csv), or None.  Not measured live: a profiler is never in a timed run."""
    for name in NCU_SUMMARIES.get((workload, kernel, specialised), []):
        if not (ROOT / "profiles" / name).exists():
            continue
        tot, unit_scale = 0.0, {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
        for line in (ROOT / "profiles" / name).read_text().splitlines():
            parts = [p.strip('"') for p in line.split('","')]
            if len(parts) == 3 and parts[0] in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
                tot += float(parts[1].replace(",", "")) * unit_scale.get(parts[2], 1.0)
        return (tot or None), f"profiles/{name}"
    return None, None


def ncu_executed_fp32(workload: str, kernel: str, specialised: bool = False):
    """Executed FP32 arithmetic of the dominant kernel from the same committed ncu capture: thread-level FADD + FMUL +
    2 x FFMA per elapsed cycle (ncu's sass_thread_inst_executed_op_f* counters) against the 2 x 128 x SMs the FFMA peak
    allows.  ncu's classes leave out FMNMX / abs / neg / MUFU, which the ALGORITHMIC count includes at one flop
    each: this is the pessimistic companion of roofline.frac, not a replacement.  -> dict or None"""
    for name in NCU_SUMMARIES.get((workload, kernel, specialised), []):
        f = ROOT / "profiles" / name
        if not f.exists():
            continue
        v = {}
        for line in f.read_text().splitlines():
            parts = [p.strip('"') for p in line.split('","')]
            if len(parts) == 3:
                v[parts[0]] = parts[1]
        try:
            fadd = float(v["smsp__sass_thread_inst_executed_op_fadd_pred_on.sum.per_cycle_elapsed"])
            fmul = float(v["smsp__sass_thread_inst_executed_op_fmul_pred_on.sum.per_cycle_elapsed"])
            ffma = float(v["smsp__sass_thread_inst_executed_op_ffma_pred_on.sum.per_cycle_elapsed"])
            peak = float(v["derived__sm__sass_thread_inst_executed_op_ffma_pred_on_x2"])
        except (KeyError, ValueError):
            return None
        def pct(key):
            try:
                return float(v[key])
            except (KeyError, ValueError):
                return None
        out = {"fadd_per_cycle": fadd, "fmul_per_cycle": fmul, "ffma_per_cycle": ffma, "flops_per_cycle": fadd + fmul + 2 * ffma,
               "peak_flops_per_cycle": peak, "frac": (fadd + fmul + 2 * ffma) / peak, "source": f"profiles/{name}",
               "fma_pipe_cycles_active_pct": pct("sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active"),
               "issue_slots_busy_pct": pct("sm__inst_issued.avg.pct_of_peak_sustained_active"),
               "note": "FADD + FMUL + 2 x FFMA thread instructions per elapsed cycle under ncu (FMNMX, abs / neg modifiers and MUFU are "
                       "not in these classes; the algorithmic count gives them one flop each)"}
        if specialised:
            out["note"] += ("; the compiled kernels do their adds / multiplies / FMAs as PACKED FADD2 / FMUL2 / FFMA2 (two rays per "
                            "instruction), which ncu's scalar op classes do not count: for them frac here covers the scalar remainder "
                            "only and fma_pipe_cycles_active_pct is the figure to read")
        return out
    return None


def cpu_sample_rows(H: int, n: int):
    n = max(1, min(n, H))
    return sorted({min(H - 1, int((i + 0.5) * H / n)) for i in range(n)})  # evenly spaced over the whole frame


def run_cpu_oracle(cmd, words, W, H, uniforms, rows, keep: bool = False):
    """Times the CPU oracle (all host threads) on single rows of the workload frame.  The oracle is test infrastructure:
    this is the cpu_baseline / --impl reference leg (and, with keep=True, the checker of the parity object), never the
    product path.  -> (Mrays/s, threads, seconds[, {row: oracle outputs incl. per-sample class / distance}])"""
    sys.path.insert(0, str(ROOT / "tests"))
    import helpers as h
    nthreads = h.oracle_lib().rmo_num_threads()
    kept = {}
    t0 = time.perf_counter()
    for r in rows:
        out = h.oracle_render(uniforms, cmd, words, W, H, row0=r, nrows=1, aov=keep)
        if keep:
            kept[r] = out
    dt = time.perf_counter() - t0
    v = len(rows) * W * 16 / dt / 1e6
    return (v, nthreads, dt, kept) if keep else (v, nthreads, dt)


def measure_cpu_baseline(cmd, words, W, H, uniforms, scene, first_rows: int = 108, target_s: float = 14.0, keep: bool = False):
    """The cpu_baseline object of the JSON line: the CPU oracle port on all host threads, on evenly spaced rows of the
    workload frame.  A first pass of `first_rows` rows sizes the sample for 10-30 s of CPU work on this box.
    keep=True also returns the oracle's rows (image, per-sample class and distance) for the parity object."""
    rows = cpu_sample_rows(H, first_rows)
    res = run_cpu_oracle(cmd, words, W, H, uniforms, rows, keep)
    if res[2] < 10.0 and len(rows) < H:
        rows = cpu_sample_rows(H, min(H, max(len(rows) + 1, int(len(rows) * target_s / max(res[2], 0.05)))))
        res = run_cpu_oracle(cmd, words, W, H, uniforms, rows, keep)
    v, nthreads, dt = res[:3]
    obj = {"value": v, "unit": "Mrays/s", "cores": nthreads, "kind": "port",
           "sample": f"{len(rows)} evenly spaced rows of the {W}x{H} {scene} frame ({len(rows) * W * 16} rays, {dt:.1f} s)"}
    return (obj, res[3]) if keep else obj


def parity_stats(img, ref_img, cls=None, ref_cls=None, t=None, ref_t=None) -> dict:
    """The three north-star gates over a set of rows: img [n, W, 4] u8, cls [n, W, 16] u8 (2 = object hit), t [n, W, 16] f32."""
    d = np.abs(img.astype(np.int32) - ref_img.astype(np.int32)).max(-1)
    out = {"rgba_gt2_frac": float((d > GATE_RGBA_LSB).mean()), "rgba_max": int(d.max()), "rgba_identical": bool(d.max() == 0)}
    if cls is not None:
        hm, hm_ref = cls == 2, ref_cls == 2
        out["hitmask_pix_frac"] = float((hm != hm_ref).any(-1).mean())
        out["hitmask_sample_frac"] = float((hm != hm_ref).mean())
        both = hm & hm_ref
        rel = np.abs(t[both] - ref_t[both]) / np.maximum(np.abs(ref_t[both]), 1e-6)
        out["t_rel_gt1e-4_frac"] = float((rel > GATE_T_REL).mean()) if rel.size else 0.0
        out["t_rel_max"] = float(rel.max()) if rel.size else 0.0
        out["t_rel_p999"] = float(np.quantile(rel, 0.999)) if rel.size else 0.0
        out["aov_identical"] = bool(np.array_equal(cls, ref_cls) and np.array_equal(t.view(np.uint32), ref_t.view(np.uint32)))
        out["gates_ok"] = bool(out["hitmask_pix_frac"] <= GATE_HITMASK and out["rgba_gt2_frac"] <= GATE_RGBA_FRAC
                               and out["t_rel_gt1e-4_frac"] <= GATE_T_FRAC)
    return out


# ---------------------------------------------------------------------------------------------- reference arm (CPU)
def reference_arm(args) -> int:
    """The reference's own CPU implementation of the path = the oracle port (the Rust/WGSL reference cannot run here).
    Loads NOTHING of the product: scenes and uniforms are the independent numpy producers of tests/helpers.py (byte-identical
    to the product's by test)."""
    if int(os.environ.get("RANK", "0")) != 0:
        return 0
    sys.path.insert(0, str(ROOT / "tests"))
    import helpers as h
    scene, W, H, n_poses, desc = WORKLOADS[args.workload]
    cmd, words = h.SCENES[scene]()
    warm = max(args.warmup, 3)
    yaw0, pitch = (0.0, 0.0) if scene == "S0" else (0.6, 0.4)
    strong = args.workload in ("cfg4", "cfg5")
    u = h.orbit_uniforms(W, H, yaw=yaw0, pitch=pitch)
    rows = cpu_sample_rows(H, 54)
    vals = []
    nthreads = 1
    for s in range(warm + args.steps):
        v, nthreads, dt = run_cpu_oracle(cmd, words, W, H, u, rows)
        if s >= warm:
            vals.append((v, dt))
        if s == 0 and dt * (warm + args.steps) > 240:  # keep the whole run within a few minutes
            rows = rows[: max(4, int(len(rows) * 240 / (dt * (warm + args.steps))))]
    v = float(np.mean([x[0] for x in vals]))
    ms = float(np.mean([x[1] for x in vals])) * 1e3
    frame_rays = W * H * 16 * (n_poses if strong else 1)
    sample = f"{len(rows)} evenly spaced rows of the {W}x{H} {scene} frame per step ({len(rows) * W * 16} rays)"
    config = {"workload": desc, "scene": scene, "width": W, "height": H, "samples_per_pixel": 16,
              "frames_per_step": n_poses if strong else max(1, args.gpus), "limits": [0.01, 100.0, 100]}
    print(json.dumps({
        "impl": "reference", "metric": "Mrays/s", "value": v, "unit": "Mrays/s", "n_gpus": args.gpus, "steps": args.steps,
        "warmup": warm, "ms_per_step": ms, "higher_is_better": True, "scaling": "strong" if strong else "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": config,
        "cpu_baseline": {"value": v, "unit": "Mrays/s", "cores": nthreads, "kind": "port", "sample": sample},
        "e2e": {"value": v, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "ms_per_step_covers": f"the bounded sample only ({len(rows)} of {H} rows), NOT the workload's whole step; value (Mrays/s) is a rate "
                              "and transfers",
        "ms_whole_step_at_this_rate": frame_rays / (v * 1e6) * 1e3,
        "note": "reference arm = CPU oracle port of ray_marching.wgsl on all host threads; the Rust/WGSL reference itself cannot run on "
                "this box (no rustc, no graphics stack: wgpu-on-lavapipe unavailable); nothing of the product is loaded by this arm"}))
    return 0


# ---------------------------------------------------------------------------------------------- B200 arm
def bind_to_gpu_cpus(index: int) -> str:
    """Multi-rank runs: keep this process — and with it the first touch of its pinned host buffers — on the CPUs NVML names
    as local to its GPU (what `numactl` would do for a launcher that does not).  The frames of cfg4 / cfg5 go back to host
    memory over N PCIe links at once, and a buffer on the far socket halves a link's rate.  Returns what it did."""
    try:
        import pynvml
        pynvml.nvmlInit()
        hnd = pynvml.nvmlDeviceGetHandleByIndex(index)
        n_cpu = os.cpu_count() or 1
        words = pynvml.nvmlDeviceGetCpuAffinity(hnd, (n_cpu + 63) // 64)
        cpus = {64 * w + b for w, bits in enumerate(words) for b in range(64) if (int(bits) >> b) & 1}
        allowed = os.sched_getaffinity(0)
        cpus &= allowed
        if not cpus:
            return "no GPU-local CPU in this process's affinity mask: unchanged"
        if cpus == allowed:
            return f"all {len(allowed)} allowed CPUs are local to GPU {index}: unchanged"
        os.sched_setaffinity(0, cpus)
        return f"bound to the {len(cpus)} CPUs local to GPU {index} (of {len(allowed)} allowed)"
    except Exception as e:  # NVML missing or no affinity information: run as launched
        return f"unchanged ({type(e).__name__}: {e})"


class Rig:
    """One rank's CUDA plumbing: the product context, a dedicated stream (the ABI reads a NULL stream as "the ctx's own
    stream", and the events that time the steps must sit on the stream the kernels are launched on), torch.distributed."""

    def __init__(self, args):
        import torch
        import rm_import
        self.torch = torch
        self.rm = rm_import.load()
        self.scenes = importlib.import_module("ray_marching_b200.scenes")
        self.sharding = importlib.import_module("ray_marching_b200.sharding")
        self.rank = int(os.environ.get("RANK", "0"))
        self.local_rank = int(os.environ.get("LOCAL_RANK", "0"))
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.cpu_binding = bind_to_gpu_cpus(self.local_rank) if self.world > 1 and args.bind_cpus == "on" else "not asked for"
        torch.cuda.set_device(self.local_rank)
        self.dev = torch.device("cuda", self.local_rank)
        self.dist = None
        if self.world > 1:
            import torch.distributed as dist
            dist.init_process_group("nccl", device_id=self.dev)
            self.dist = dist
        rm = self.rm
        self.ctx = rm.Context(self.local_rank)
        self.ctx.set_mode(rm.MODE_FAST if args.mode == "fast" else rm.MODE_EXACT)
        self.ctx.set_kernel({"wavefront": rm.KERNEL_WAVEFRONT, "simple": rm.KERNEL_SIMPLE, "wavefront2": rm.KERNEL_WAVEFRONT2,
                             "direct": rm.KERNEL_DIRECT}[args.kernel])
        if args.early_out:
            self.ctx.set_early_out(True)
        if args.cull == "on":
            self.ctx.set_cull(True)
        self.stream = torch.cuda.Stream(self.dev)
        torch.cuda.set_stream(self.stream)
        self.sp = self.stream.cuda_stream
        assert self.sp != 0
        self.flush = torch.empty(256 << 20, dtype=torch.uint8, device=self.dev)

    def barrier(self):
        if self.dist is not None:
            self.dist.barrier()
        self.torch.cuda.synchronize(self.dev)
        self.ctx.sync()

    def time_steps(self, step, n: int, mid=None):
        """n timed steps, L2 flushed (untimed) before each; CUDA events on the launching stream.  -> per-step ms list.
        mid: optional callable run first inside a step with an event behind it (per-rank paint time) -> (ms, mid ms)."""
        torch = self.torch
        ev0 = [torch.cuda.Event(enable_timing=True) for _ in range(n)]
        ev1 = [torch.cuda.Event(enable_timing=True) for _ in range(n)]
        evm = [torch.cuda.Event(enable_timing=True) for _ in range(n)] if mid else None
        for i in range(n):
            self.flush.zero_()            # L2 flush, outside the timed event pair
            ev0[i].record(self.stream)
            if mid:
                mid()
                evm[i].record(self.stream)
            step()
            ev1[i].record(self.stream)
        self.barrier()
        ms = [a.elapsed_time(b) for a, b in zip(ev0, ev1)]
        if mid:
            return ms, [a.elapsed_time(b) for a, b in zip(ev0, evm)]
        return ms

    def max_over_ranks(self, *vals):
        if self.dist is None:
            return [float(v) for v in vals]
        tt = self.torch.tensor([float(v) for v in vals], device=self.dev, dtype=self.torch.float64)
        self.dist.all_reduce(tt, op=self.dist.ReduceOp.MAX)
        return [float(x) for x in tt]

    def gather_floats(self, v: float):
        if self.dist is None:
            return [float(v)]
        tt = self.torch.zeros(self.world, device=self.dev, dtype=self.torch.float64)
        tt[self.rank] = float(v)
        self.dist.all_reduce(tt)
        return [float(x) for x in tt]

    def specialise(self, args) -> dict:
        """Program specialisation of the prepared command stream (once, untimed; cached process-wide)."""
        rm, ctx = self.rm, self.ctx
        spec = {"requested": args.specialize == "on" and args.kernel == "direct" and not args.early_out and args.cull != "on", "active": False}
        if spec["requested"]:
            try:
                ctx.specialize()
            except rm.RmError as e:  # no NVRTC on the box, or the program is too long: the interpreter kernel runs, and the line says so
                spec["reason"] = str(e)[:300]
            spec.update({k: v for k, v in ctx.spec_info().items() if k in ("active", "rays_per_thread", "min_blocks", "compile_ms", "cubin_bytes")})
            spec["tier"] = {0: "interpreter", 1: "structure (parameters as kernel arguments)", 2: "literal (parameters as immediates)"}[ctx.spec_tier]
        return spec

    def counters_of(self, count_step):
        self.ctx.cull_stats()
        self.ctx.set_counters(True)
        count_step()
        c = self.ctx.get_counters()
        self.ctx.set_counters(False)
        self.cull_stats = self.ctx.cull_stats()   # executed map_scene flops of that step when culling is on
        return c


def kernel_label(args, spec) -> str:
    if spec.get("active"):
        return f"direct+specialised(rm_spec,R={spec.get('rays_per_thread')},{spec.get('tier', '?').split(' ')[0]} tier)"
    return args.kernel + ("+interpreter" if args.kernel == "direct" else "")


class SharedHostFrame:
    """One host buffer all ranks of the box can DMA into: POSIX shared memory, page-locked in every process
    (rm_host_register).  With one rank it is a plain pinned buffer."""

    def __init__(self, rig: Rig, nbytes: int, tag: str):
        self.rig, self.nbytes = rig, nbytes
        self.shm = None
        self.pinned = None
        if rig.dist is None:
            self.pinned = rig.rm.PinnedBuffer(nbytes)
            self.ptr, self.array = self.pinned.ptr, self.pinned.array
            return
        from multiprocessing import shared_memory
        name = f"rm_b200_{tag}_{os.environ.get('MASTER_PORT', '0')}"
        if rig.rank == 0:
            try:
                old = shared_memory.SharedMemory(name=name)
                old.close()
                old.unlink()
            except Exception:
                pass
            self.shm = shared_memory.SharedMemory(name=name, create=True, size=nbytes)
        rig.barrier()
        if rig.rank != 0:
            self.shm = shared_memory.SharedMemory(name=name)
        self.array = np.ndarray((nbytes,), dtype=np.uint8, buffer=self.shm.buf)
        self.ptr = self.array.ctypes.data
        rig.rm.host_register(self.ptr, nbytes)
        rig.barrier()

    def close(self):
        rig = self.rig
        if self.pinned is not None:
            self.pinned.free()
            return
        rig.rm.host_unregister(self.ptr)
        self.array = None
        rig.barrier()
        try:
            self.shm.close()
            if rig.rank == 0:
                self.shm.unlink()
            else:  # the resource tracker of an attaching process would otherwise unlink the segment again at exit
                from multiprocessing import resource_tracker
                resource_tracker.unregister(self.shm._name, "shared_memory")
        except Exception:
            pass


# ---- cfg4: one 8K frame, interleaved strips --------------------------------------------------------------------------
def run_cfg4(rig: Rig, args, steps: int, warm: int, with_n1: bool) -> dict:
    rm, ctx, torch, dist = rig.rm, rig.ctx, rig.torch, rig.dist
    rank, world, dev, sp = rig.rank, rig.world, rig.dev, rig.sp
    scene, W, H, _, desc = WORKLOADS["cfg4"]
    cmd, words = rig.scenes.program(scene)
    uniforms = rig.scenes.orbit_uniforms(W, H, yaw=0.6, pitch=0.4)
    strip = rig.sharding.balanced_strip_rows(H, world)   # every rank the same number of strips (16 rows when that works)
    ctx.prepare(uniforms, cmd, words)
    spec = rig.specialise(args)
    out = {"workload": desc, "ranks": world, "kernel": kernel_label(args, spec)}
    rows_local = rig.sharding.local_rows(H, strip, rank, world)
    rays_total = W * H * 16

    # rank 0 alone, the whole frame (same harness): the N=1 time of THIS box
    n1_ms = None
    if with_n1:
        if rank == 0:
            for _ in range(max(1, warm - 1)):
                ctx.paint(W, H, 0, H, None, 0, sp)
            torch.cuda.synchronize(dev)
            evs = []
            for _ in range(max(2, min(steps, 4))):
                rig.flush.zero_()
                a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a.record(rig.stream)
                ctx.paint(W, H, 0, H, None, 0, sp)
                b.record(rig.stream)
                evs.append((a, b))
            torch.cuda.synchronize(dev)
            n1_ms = float(np.mean([a.elapsed_time(b) for a, b in evs]))
        rig.barrier()

    band = None
    frame_ptr = None
    frame_out = {}
    p2p = dist is not None and args.gather == "p2p"
    if dist is None:
        # one GPU: the "shard" is every strip, painted into the ctx-owned frame
        def paint():
            ctx.paint_strips(W, H, strip, 0, 1, None, 0, sp)

        def finish():
            pass
        out["gather"] = "none (one rank paints every strip)"
        payload = 0
    elif p2p:
        # rank 0 owns the frame; the others map it (CUDA IPC) and their render kernels store into it over NVLink
        hbuf = torch.zeros(64, dtype=torch.uint8, device=dev)
        if rank == 0:
            frame_ptr, hnd = ctx.shared_frame_create(W * H * 4)
            hbuf.copy_(torch.frombuffer(bytearray(hnd), dtype=torch.uint8))
        dist.broadcast(hbuf, src=0)
        if rank != 0:
            frame_ptr = ctx.shared_frame_open(bytes(hbuf.cpu().numpy().tobytes()))
        done = torch.zeros(1, device=dev)

        def paint():
            ctx.paint_strips_into(W, H, strip, rank, world, frame_ptr, W * 4, sp)

        def finish():
            dist.all_reduce(done)  # enqueued behind the paint on every rank: rank 0 passes it once all strips are stored
        out["gather"] = ("every rank's render kernel stores its strips straight into rank 0's frame over NVLink (CUDA IPC mapping); "
                         "completion = one 4-byte all-reduce")
        payload = (H - rig.sharding.local_rows(H, strip, 0, world)) * W * 4
    else:
        rows_max = rig.sharding.max_local_rows(H, strip, world)
        band = torch.zeros((rows_max, W, 4), dtype=torch.uint8, device=dev)
        gathered = [torch.zeros_like(band) for _ in range(world)] if rank == 0 else None

        def paint():
            ctx.paint_strips(W, H, strip, rank, world, band.data_ptr(), W * 4, sp)

        def finish():
            dist.gather(band, gathered, dst=0)
            if rank == 0:
                frame_out["f"] = rig.sharding.reassemble(torch.stack(gathered), H, strip, world)
        out["gather"] = "packed bands + NCCL gather + indexed re-order on rank 0"
        payload = (world - 1) * rows_max * W * 4

    for _ in range(warm):
        paint()
        finish()
    rig.barrier()
    ms_list, paint_list = rig.time_steps(finish, steps, mid=paint)
    ms_local, paint_local = float(np.mean(ms_list)), float(np.mean(paint_list))
    (ms,) = rig.max_over_ranks(ms_local)
    paint_per_rank = rig.gather_floats(paint_local)
    out.update(ms=ms, mrays=rays_total / (ms * 1e-3) / 1e6, fps=1e3 / ms, steps=steps, strip_rows=strip,
               paint_ms_per_rank=paint_per_rank,
               paint_imbalance=max(paint_per_rank) / (sum(paint_per_rank) / len(paint_per_rank)) - 1.0,
               completion_ms=ms - max(paint_per_rank),
               bytes_per_step={"nvlink_p2p_or_nccl_payload": int(payload), "completion_allreduce": 4 if p2p else 0})
    if n1_ms is None and dist is None:
        n1_ms = ms
    (n1_all,) = rig.max_over_ranks(n1_ms or 0.0)
    out["n1_ms_same_box"] = n1_all or None
    if n1_all:
        out["speedup_vs_n1_same_box"] = n1_all / ms

    # ---- e2e: host uniforms + words in, the whole frame back in (pinned) host memory.
    # N > 1: every rank copies ITS OWN strips from its device band to their rows of one shared pinned host frame, so the
    # frame leaves over N PCIe links at once (the device-resident frame on GPU 0 above is for device-side consumers)
    n_e2e = max(3, min(steps, 5))
    host = SharedHostFrame(rig, W * H * 4, tag="cfg4")
    if dist is not None and band is None:
        band = torch.zeros((rig.sharding.max_local_rows(H, strip, world), W, 4), dtype=torch.uint8, device=dev)
    for it in range(1 + n_e2e):
        if it == 1:
            rig.barrier()
            t0 = time.perf_counter()
            up0 = ctx.program_uploads
        ctx.prepare(uniforms, cmd, words)
        if dist is None:
            ctx.paint(W, H, 0, H, None, 0, sp)
            ctx.copy_strips_to_host(None, W, H, strip, 0, 1, host.ptr, sp)
        else:
            ctx.paint_strips(W, H, strip, rank, world, band.data_ptr(), W * 4, sp)
            ctx.copy_strips_to_host(band.data_ptr(), W, H, strip, rank, world, host.ptr, sp)
        rig.barrier()   # every rank's strips are in the host frame
    e2e_local = (time.perf_counter() - t0) * 1e3 / n_e2e
    (e2e_ms,) = rig.max_over_ranks(e2e_local)
    out["e2e_ms"] = e2e_ms
    out["e2e_mrays"] = rays_total / (e2e_ms * 1e-3) / 1e6
    out["e2e_mode"] = ("each rank paints its strips into its own device band and copies them to their rows of ONE shared pinned host frame "
                       f"({world} PCIe links in parallel, one strided cudaMemcpy2DAsync per rank)") if dist is not None else \
        "rm_paint + one copy of the frame to pinned host memory"
    out["h2d_bytes_per_step"] = 144 + 12 + (ctx.program_uploads - up0) * (4 + int(len(words)) * 4) // n_e2e
    out["d2h_bytes_per_step"] = W * H * 4
    # untimed checks: both assembled frames (device frame on rank 0, shared host frame) must equal rank 0's own full render
    if rank == 0:
        alone = np.empty((H, W, 4), np.uint8)
        ctx.paint(W, H)
        ctx.download(alone)
        host_ok = bool(np.array_equal(alone.reshape(-1), host.array))
        if p2p:
            devf = np.empty((H, W, 4), np.uint8)
            ctx.copy_to_host(frame_ptr, devf, W * H * 4, sp)
            dev_ok = bool(np.array_equal(alone, devf))
        elif dist is not None:
            dev_ok = bool(np.array_equal(alone, frame_out["f"][:H].cpu().numpy()))
        else:
            dev_ok = True
        out["frame_check"] = dev_ok and host_ok
        out["frame_check_detail"] = {"device_frame_on_rank0": dev_ok, "shared_host_frame": host_ok}
    rig.barrier()
    host.close()
    if p2p and rank != 0:
        ctx.shared_frame_close(frame_ptr)
    rig.barrier()
    if p2p and rank == 0:
        ctx.shared_frame_close(frame_ptr)
    out["_spec"] = spec
    out["_local"] = {"ms_local": paint_local if dist is not None else ms_local, "uniforms": uniforms, "cmd": cmd, "words": words,
                     "count_step": (lambda: ctx.paint_strips(W, H, strip, rank, world, None, 0, sp)), "n_pixels_local": rows_local * W}
    return out


# ---- cfg5: 1024 orbit poses of S1 at 1080p ----------------------------------------------------------------------------
def run_cfg5(rig: Rig, args, steps: int, warm: int, with_n1: bool, n_poses: int = 1024) -> dict:
    rm, ctx, torch, dist = rig.rm, rig.ctx, rig.torch, rig.dist
    rank, world, dev, sp = rig.rank, rig.world, rig.dev, rig.sp
    scene, W, H, _, desc = WORKLOADS["cfg5"]
    cmd, words = rig.scenes.program(scene)
    all_u = rig.scenes.orbit_pose_batch(W, H, n_poses, pitch=0.4)
    my = list(rig.sharding.pose_slice(n_poses, rank, world))
    my_u = b"".join(all_u[k * 144:(k + 1) * 144] for k in my)
    ctx.prepare(all_u[:144], cmd, words)
    spec = rig.specialise(args)
    out = {"workload": desc, "ranks": world, "poses": n_poses, "kernel": kernel_label(args, spec),
           "sharding": "pose k -> rank k % N; frames stay on their GPU (throughput mode); no data-path collective"}
    rays_total = n_poses * W * H * 16

    n1_ms = None
    if with_n1:
        if rank == 0:
            ctx.paint_poses(all_u, min(n_poses, 64), W, H, None, 0, 0, sp)  # warm-up
            torch.cuda.synchronize(dev)
            evs = []
            for _ in range(2):
                rig.flush.zero_()
                a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a.record(rig.stream)
                ctx.paint_poses(all_u, n_poses, W, H, None, 0, 0, sp)
                b.record(rig.stream)
                evs.append((a, b))
            torch.cuda.synchronize(dev)
            n1_ms = float(np.mean([a.elapsed_time(b) for a, b in evs]))
        rig.barrier()

    def step():
        ctx.paint_poses(my_u, len(my), W, H, None, 0, 0, sp)
    for _ in range(max(1, min(warm, 2))):
        step()
    rig.barrier()
    ms_list = rig.time_steps(step, steps)
    ms_local = float(np.mean(ms_list))
    (ms,) = rig.max_over_ranks(ms_local)
    per_rank = rig.gather_floats(ms_local)
    out.update(ms=ms, mrays=rays_total / (ms * 1e-3) / 1e6, fps=n_poses * 1e3 / ms, steps=steps, ms_per_rank=per_rank,
               imbalance=max(per_rank) / (sum(per_rank) / len(per_rank)) - 1.0,
               bytes_per_step={"nvlink_p2p_or_nccl_payload": 0, "completion_allreduce": 0})
    if n1_ms is None and dist is None:
        n1_ms = ms
    (n1_all,) = rig.max_over_ranks(n1_ms or 0.0)
    out["n1_ms_same_box"] = n1_all or None
    if n1_all:
        out["speedup_vs_n1_same_box"] = n1_all / ms

    # e2e: host uniform blocks in, EVERY frame back in host memory (rm_paint_poses_host: the D2H copy of frame p overlaps
    # the render of frame p+1); the frames land in a pinned 32-frame host ring, as a consumer would drain them
    ring_frames = 32
    pinned = rm.PinnedBuffer(ring_frames * W * H * 4)
    n_e2e = 2
    for it in range(1 + n_e2e):
        if it == 1:
            rig.barrier()
            t0 = time.perf_counter()
            up0 = ctx.program_uploads
        ctx.prepare(all_u[:144], cmd, words)
        ctx.paint_poses_host(my_u, len(my), W, H, pinned.ptr, W * H * 4, ring_frames)
        rig.barrier()
    e2e_local = (time.perf_counter() - t0) * 1e3 / n_e2e
    (e2e_ms,) = rig.max_over_ranks(e2e_local)
    out["e2e_ms"] = e2e_ms
    out["e2e_mrays"] = rays_total / (e2e_ms * 1e-3) / 1e6
    out["h2d_bytes_per_step"] = n_poses * 144 + 12 + (ctx.program_uploads - up0) * (4 + int(len(words)) * 4) // n_e2e
    out["d2h_bytes_per_step"] = n_poses * W * H * 4
    # untimed check: the last pose of this rank's batch (left in the host ring) equals the same pose rendered alone
    last = (len(my) - 1) % ring_frames
    got = pinned.array[last * W * H * 4:(last + 1) * W * H * 4].reshape(H, W, 4).copy()
    alone = ctx.render(my_u[-144:], cmd, words, W, H)
    (bad,) = rig.max_over_ranks(0.0 if np.array_equal(got, alone) else 1.0)
    out["frame_check"] = bad == 0.0
    pinned.free()
    out["_spec"] = spec
    out["_local"] = {"ms_local": ms_local, "uniforms": all_u[:144], "cmd": cmd, "words": words, "count_step": step,
                     "n_pixels_local": len(my) * W * H}
    return out


def public(d: dict) -> dict:
    return {k: v for k, v in d.items() if not k.startswith("_")}


def parity_object(rig: Rig, args, W, H, uniforms, cmd, words, oracle_rows: dict, timed_frame, spec) -> dict:
    """Oracle-vs-CUDA parity on the benchmarked frame at its full size: the rows the cpu_baseline leg rendered with the
    oracle against the same rows of (a) the frame the TIMED kernel produced, (b) that kernel's own per-sample classes and
    hit distances (rm_paint_aov: the production kernel body with its AOVs written out), for the timed kernel and for the
    interpreter, in the benchmarked arithmetic mode; plus exact mode, which must be identical bit for bit."""
    rm, ctx = rig.rm, rig.ctx
    rows = sorted(oracle_rows)
    ref_img = np.stack([oracle_rows[r]["rgba"][0] for r in rows])
    ref_cls = np.stack([oracle_rows[r]["cls"][0] for r in rows])
    ref_t = np.stack([oracle_rows[r]["t"][0] for r in rows])

    def aov_rows():
        img, cls, t, flags = [], [], [], []
        for r in rows:
            a = ctx.paint_aov(W, H, r, 1)
            img.append(a["rgba"][0])
            cls.append(a["cls"][0])
            t.append(a["t"][0])
            flags.append(a["specialised"])
        return np.stack(img), np.stack(cls), np.stack(t), all(flags)

    out = {"rows": len(rows), "size": f"{W}x{H}", "mode": args.mode,
           "oracle": "oracle/rm_oracle.c (CPU port of ray_marching.wgsl; parity unpinned by the reference, which holds no vectors)",
           "gates": {"hitmask_pix_frac": GATE_HITMASK, "rgba_gt2_frac": GATE_RGBA_FRAC, "t_rel": GATE_T_REL,
                     "t_rel_note": "the north star states no percentile; applied as <= 1e-4 on >= 99.9 % of the samples hit by both (a flipped "
                                   "d < min_dist moves one sample by one march step); the observed max and fraction are printed with it"}}
    ctx.prepare(uniforms, cmd, words)
    legs = ([("timed_kernel", True), ("structure_tier", True)] if spec.get("active") else []) + [("interpreter", False)]

    def select(name, use_spec):
        ctx.set_spec_literal_after(0 if name == "structure_tier" else 8)
        if use_spec:
            ctx.specialize()
        else:
            ctx.set_specialize(False)

    for name, use_spec in legs:
        select(name, use_spec)
        img, cls, t, was_spec = aov_rows()
        st = parity_stats(img, ref_img, cls, ref_cls, t, ref_t)
        st["kernel"] = (kernel_label(args, spec) if name == "timed_kernel" else "direct+specialised(structure tier)") if use_spec else "direct+interpreter"
        st["aov_run_specialised"] = bool(was_spec)
        if timed_frame is not None and (name == "timed_kernel" or not spec.get("active")):
            tf = timed_frame[rows]
            st["timed_frame_rows"] = parity_stats(tf, ref_img)
            st["aov_rgba_equals_timed_frame_rows"] = bool(np.array_equal(tf, img))
        out[name] = st
    if not spec.get("active"):
        out["timed_kernel"] = out["interpreter"]
    # flat copies of the timed kernel's numbers (the keys VERDICT r1 asked for)
    for k in ("hitmask_pix_frac", "rgba_gt2_frac", "rgba_max", "t_rel_gt1e-4_frac", "t_rel_max", "gates_ok"):
        out[k] = out["timed_kernel"][k]
    # exact mode: bit-identical rows (image, classes, distance bit patterns) for both kernels
    mode0 = rm.MODE_FAST if args.mode == "fast" else rm.MODE_EXACT
    ctx.set_mode(rm.MODE_EXACT)
    ident = {}
    for name, use_spec in legs:
        try:
            select(name, use_spec)
            img, cls, t, _ = aov_rows()
            ident[name] = bool(np.array_equal(img, ref_img) and np.array_equal(cls, ref_cls)
                               and np.array_equal(t.view(np.uint32), ref_t.view(np.uint32)))
        except rm.RmError as e:
            ident[name] = f"not run: {str(e)[:120]}"
    out["exact_mode_rows_identical"] = bool(all(v is True for v in ident.values()))
    out["exact_mode_rows_identical_detail"] = ident
    ctx.set_mode(mode0)
    ctx.set_spec_literal_after(8)
    ctx.set_specialize(False)
    if spec.get("active"):
        ctx.specialize()
    return out


def b200_arm(args) -> int:
    rig = Rig(args)
    rm, ctx, dist = rig.rm, rig.ctx, rig.dist
    rank, world, sp = rig.rank, rig.world, rig.sp
    scene, W, H, n_poses, desc = WORKLOADS[args.workload]
    if args.workload == "cfg5" and args.poses:
        n_poses = args.poses
    warm = max(args.warmup, 3)
    yaw0, pitch = (0.0, 0.0) if scene == "S0" else (0.6, 0.4)
    strong = args.workload in ("cfg4", "cfg5")
    config = {"workload": desc, "scene": scene, "width": W, "height": H, "samples_per_pixel": 16,
              "frames_per_step": n_poses if strong else world, "limits": [0.01, 100.0, 100],
              "l2": "flushed between timed steps (256 MiB memset, untimed)"}
    if args.early_out:
        config["early_out"] = "ON: executed work per map_scene evaluation is less than the algorithmic FLOPs the roofline fraction counts"
    if args.cull == "on":
        config["cull"] = ("ON (opt-in, never the headline): every block interprets only the instructions its tile's cone can be changed by "
                          "(csrc/rm_cull.cu); frames and work counters are identical to the unculled paint, the work per evaluation is not "
                          "the shader's literal arithmetic: roofline.frac counts ALGORITHMIC flops and overstates, frac_executed counts the "
                          "flops of the instructions actually run")
    if world > 1:
        config["cpu_binding_rank0"] = rig.cpu_binding
    peaks = measured_peaks()
    hbm_gbs = float(peaks.get("hbm_gbs", 0.0)) or None
    extra = {}
    interp = parity = cpu_base = None

    if strong:
        # the sharded config is the headline of this line
        with ClockSampler(rig.local_rank) as clk:
            t_wall0 = time.perf_counter()
            launches0 = ctx.launch_count
            res = run_cfg4(rig, args, args.steps, warm, with_n1=world > 1) if args.workload == "cfg4" else \
                run_cfg5(rig, args, max(2, min(args.steps, 6)), warm, with_n1=world > 1, n_poses=n_poses)
            t_wall = time.perf_counter() - t_wall0
            launches = ctx.launch_count - launches0
        loc, spec = res["_local"], res["_spec"]
        cmd, words, uniforms = loc["cmd"], loc["words"], loc["uniforms"]
        ctx.prepare(uniforms, cmd, words)
        counters = rig.counters_of(loc["count_step"])
        n_pixels_local = loc["n_pixels_local"]
        ms, ms_local, value = res["ms"], loc["ms_local"], res["mrays"]
        rays_step_total = (n_poses if args.workload == "cfg5" else 1) * W * H * 16
        e2e = {"value": res["e2e_mrays"], "unit": "Mrays/s", "ms_per_step": res["e2e_ms"], "h2d_bytes_per_step": res["h2d_bytes_per_step"],
               "d2h_bytes_per_step": res["d2h_bytes_per_step"]}
        config.update(sharding=res.get("gather") or res.get("sharding"), frame_check=res.get("frame_check"))
        extra["strong"] = {args.workload: public(res)}
    else:
        # ---- single frame per rank per step (cfg3 headline; cfg1 / cfg2 / n4)
        cmd, words = rig.scenes.program(scene)  # flattened by the C++ host (BuildCommands)
        # weak scaling = FIXED work per GPU: every rank renders the same configs[2] pose.  (A different yaw per rank changes
        # the ray-march work per rank; measured on 8 GPUs the costliest of 8 yaws took 47.19 ms against 44.01 ms for this
        # pose, which is view imbalance, not scaling loss — profiles/r1_s13_NOTES.md.)
        uniforms = rig.scenes.orbit_uniforms(W, H, yaw=yaw0, pitch=pitch)
        ctx.prepare(uniforms, cmd, words)
        rays_step_total = W * H * 16 * world
        n_pixels_local = W * H
        config.update(sharding="one frame per rank per step (the same pose on every rank: fixed work per GPU), no collective")

        def step():
            ctx.paint(W, H, 0, H, None, 0, sp)

        spec = rig.specialise(args)
        if spec["requested"]:
            spec["note"] = ("map_scene of this command stream compiled into the kernel at run time (NVRTC) outside the timed region, like the "
                            "reference's shader compile at start-up.  Two tiers: one kernel per program STRUCTURE with the values as kernel "
                            "arguments (an edit of a value costs no compile), and, for words that stay put — this static scene — one with the "
                            "values as immediates (the headline; value_structure_tier is the other).  Same arithmetic per ray, exact mode "
                            "bit-identical to the oracle")
        counters = rig.counters_of(step)  # work counters of this rank's step (untimed) -> algorithmic FLOPs

        for _ in range(warm):
            step()
        rig.barrier()
        launches0 = ctx.launch_count
        spec_launches0 = ctx.spec_info()["launches"]
        with ClockSampler(rig.local_rank) as clk:
            t_wall0 = time.perf_counter()
            per_step = rig.time_steps(step, args.steps)
            t_wall = time.perf_counter() - t_wall0
        launches = ctx.launch_count - launches0
        ms_local = float(np.mean(per_step))
        timed_frame = None
        if rank == 0 and world == 1 and args.parity == "on" and not args.no_cpu_baseline:
            timed_frame = np.empty((H, W, 4), np.uint8)   # the frame the timed kernel left in the ctx-owned buffer
            ctx.download(timed_frame)
        if spec["active"]:
            spec["launches_in_timed_region"] = ctx.spec_info()["launches"] - spec_launches0
            # the same steps through the interpreter kernel (what paints until a program has been compiled)
            ctx.set_specialize(False)
            step()
            rig.barrier()
            n_i = max(3, min(args.steps, 10))
            interp_ms_local = float(np.mean(rig.time_steps(step, n_i)))
            (interp_ms,) = rig.max_over_ranks(interp_ms_local)
            interp = {"ms": interp_ms, "ms_local": interp_ms_local, "steps": n_i}
            spec["interpreter_ms_per_step_this_rank"] = interp_ms_local
            ctx.specialize()  # back on (a cache hit) for the end-to-end leg
            # and through the structure-tier kernel: what paints right after an edit of a value
            ctx.set_spec_literal_after(0)
            step()
            rig.barrier()
            if ctx.spec_tier == 1:
                st_ms_local = float(np.mean(rig.time_steps(step, n_i)))
                (st_ms,) = rig.max_over_ranks(st_ms_local)
                extra["_structure"] = {"ms": st_ms, "ms_local": st_ms_local}
            ctx.set_spec_literal_after(8)
        elif args.kernel == "direct" and not args.early_out and args.cull != "on":
            interp = {"ms": None, "ms_local": ms_local, "steps": args.steps}

        # ---- e2e: host buffers in, host result out, through the public C-ABI calls (prepare + paint + download)
        n_e2e = max(3, min(args.steps, 10))

        def e2e_pipelined():
            ring_frames = 3
            pinned = rm.PinnedBuffer(ring_frames * W * H * 4)
            batch = uniforms * n_e2e                   # one 144-byte block per step, in host memory
            ctx.prepare(uniforms, cmd, words)
            ctx.paint_poses_host(batch[:2 * 144], 2, W, H, pinned.ptr, W * H * 4, ring_frames)  # warm-up (allocates the device ring)
            rig.barrier()
            t0 = time.perf_counter()
            up0 = ctx.program_uploads
            ctx.prepare(uniforms, cmd, words)          # unchanged words: only the uniforms are replaced
            ctx.paint_poses_host(batch, n_e2e, W, H, pinned.ptr, W * H * 4, ring_frames)  # returns with every frame in host memory
            ms_ = (time.perf_counter() - t0) * 1e3 / n_e2e
            ups = ctx.program_uploads - up0
            pinned.free()
            return ms_, ups

        def e2e_serial():
            pinned = rm.PinnedBuffer(W * H * 4)
            for it in range(2 + n_e2e):
                if it == 2:
                    rig.barrier()
                    t0 = time.perf_counter()
                    up0 = ctx.program_uploads
                ctx.prepare(uniforms, cmd, words)      # H2D: uniforms + (when changed) the decoded program
                ctx.paint(W, H)
                ctx.download(pinned)                   # D2H: the RGBA8 frame (synchronises)
            ms_ = (time.perf_counter() - t0) * 1e3 / n_e2e
            ups = ctx.program_uploads - up0
            pinned.free()
            return ms_, ups

        PIPE_MODE = "pipelined: rm_paint_poses_host, frame k copied to host while frame k+1 renders"
        other_ms_local = None
        if args.e2e == "pipelined":
            e2e_ms_local, uploads = e2e_pipelined()
            config["e2e_mode"] = PIPE_MODE
        else:
            e2e_ms_local, uploads = e2e_serial()
            other_ms_local, _ = e2e_pipelined()        # reported beside the headline (e2e.pipelined), never instead of it
        # uniforms + limits travel with every launch; the program only when rm_prepare really uploaded it
        h2d, d2h = 144 + 12 + uploads * (4 + int(len(words)) * 4) // n_e2e, W * H * 4
        ms, e2e_ms = rig.max_over_ranks(ms_local, e2e_ms_local)
        value = rays_step_total / (ms * 1e-3) / 1e6
        e2e = {"value": rays_step_total / (e2e_ms * 1e-3) / 1e6, "unit": "Mrays/s", "ms_per_step": e2e_ms,
               "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h}
        if other_ms_local is not None:
            (other_ms,) = rig.max_over_ranks(other_ms_local)
            e2e["pipelined"] = {"value": rays_step_total / (other_ms * 1e-3) / 1e6, "ms_per_step": other_ms, "mode": PIPE_MODE,
                                "note": "the same steps with the copy of frame k overlapping the paint of frame k+1 (what --e2e pipelined "
                                        "makes the headline); the e2e value above is the serial one: every step waits for its frame"}

        # ---- cpu_baseline (rank 0, N=1) and the parity object: the oracle's rows against the same rows of the timed frame
        if rank == 0 and world == 1 and not args.no_cpu_baseline:
            if args.parity == "on":
                cpu_base, oracle_rows = measure_cpu_baseline(cmd, words, W, H, uniforms, scene, keep=True)
                parity = parity_object(rig, args, W, H, uniforms, cmd, words, oracle_rows, timed_frame, spec)
            else:
                cpu_base = measure_cpu_baseline(cmd, words, W, H, uniforms, scene)

        # ---- the sharded configs of the north star, in the same invocation (default workload only)
        if args.workload == "cfg3" and args.strong == "on":
            extra["strong"] = {
                "cfg4": public(run_cfg4(rig, args, max(3, min(args.steps, 5)), 2, with_n1=world > 1)),
                "cfg5": public(run_cfg5(rig, args, max(2, min(args.steps, 4)), 1, with_n1=world > 1, n_poses=args.poses or 1024)),
                "note": "strong scaling of the two sharded north-star configs, measured in this same run; n1_ms_same_box = rank 0 doing the "
                        "whole step alone on this box before the group step (at N = 1: the step itself)"}

    if rank == 0:
        fmap = f_map(words, cmd)
        flops = frame_flops(counters, fmap, n_pixels_local)
        peak = ctx.fp32_peak()
        achieved = flops / (ms_local * 1e-3) / 1e12  # rank 0's kernel work over rank 0's device time
        fb_bytes = n_pixels_local * 4
        traffic, traffic_src = ncu_traffic(args.workload, args.kernel, spec["active"]) if args.mode == "fast" else (None, None)
        roof = {"bound": "fp32", "achieved": achieved, "peak": peak["tflops"], "unit": "TFLOP/s", "frac": achieved / peak["tflops"],
                "frac_of_nominal_74.4": achieved / 74.4, "kernel": kernel_label(args, spec),
                "traffic": traffic, "traffic_unit": "bytes of DRAM read+write per launch", "traffic_source": traffic_src,
                "algorithmic_bytes_per_launch": fb_bytes + 144 + 12 + 4 + int(len(words)) * 4,
                "peak_source": "FFMA micro-benchmark run live on this GPU (rm_fp32_peak); MEASURED_PEAKS.json has no FP32 ALU peak; "
                               "nominal 148 SM x 128 lanes x 2 x 1.965 GHz = 74.4",
                "algorithmic_flops_per_step": flops, "f_map_flops_per_eval": fmap,
                "counters": {"n_step": int(counters[0]), "n_hit": int(counters[1]), "n_floor": int(counters[2]), "n_sample": int(counters[3])},
                "framebuffer_gbs": fb_bytes / (ms_local * 1e-3) / 1e9,
                "hbm_peak_gbs": hbm_gbs, "hbm_peak_source": "MEASURED_PEAKS.json hbm_gbs" if hbm_gbs else "MEASURED_PEAKS.json absent",
                "framebuffer_frac_of_hbm_peak": (fb_bytes / (ms_local * 1e-3) / 1e9 / hbm_gbs) if hbm_gbs else None}
        if args.mode == "fast" and args.cull != "on":
            roof["executed_fp32_ncu"] = ncu_executed_fp32(args.workload, args.kernel, spec["active"])
        if args.cull == "on" and getattr(rig, "cull_stats", {}).get("culled_paints"):
            n_eval = float(counters[0]) + 4.0 * float(counters[1])
            executed = flops - n_eval * fmap + float(rig.cull_stats["executed_map_flops"])
            roof.update(executed_flops_per_step=executed, frac_executed=executed / (ms_local * 1e-3) / 1e12 / peak["tflops"],
                        executed_map_flops_per_eval=float(rig.cull_stats["executed_map_flops"]) / max(n_eval, 1.0),
                        cull_note="frac counts the algorithmic flops of the FULL program per evaluation (overstates); frac_executed counts "
                                  "the instructions the tiles kept; the culling pre-pass (one thread per tile) is inside the timed step")
        if interp:
            roof["achieved_interpreter"] = flops / (interp["ms_local"] * 1e-3) / 1e12
            roof["frac_interpreter"] = roof["achieved_interpreter"] / peak["tflops"]
        structure = extra.pop("_structure", None)
        if structure:
            roof["frac_structure_tier"] = flops / (structure["ms_local"] * 1e-3) / 1e12 / peak["tflops"]
        line = {"metric": "Mrays/s", "value": value, "unit": "Mrays/s", "n_gpus": world, "steps": args.steps, "warmup": warm,
                "ms_per_step": ms, "fps": (n_poses if strong else world) * 1e3 / ms, "higher_is_better": True,
                "scaling": "strong" if strong else "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "config": dict(config, mode=args.mode, kernel=kernel_label(args, spec), specialize=spec), "clocks": clk.summary(),
                "gpu_launches": int(launches), "roofline": roof, "wall_s_timed_region": t_wall}
        if interp:
            ims = interp["ms"] if interp.get("ms") else ms
            line["value_interpreter"] = rays_step_total / (ims * 1e-3) / 1e6
            line["ms_per_step_interpreter"] = ims
            line["interpreter_note"] = ("the decoded-program interpreter kernel (k_direct) over the same steps in the same harness: what paints "
                                        "until a program has been compiled")
        if structure:
            line["value_structure_tier"] = rays_step_total / (structure["ms"] * 1e-3) / 1e6
            line["ms_per_step_structure_tier"] = structure["ms"]
        line["e2e"] = e2e
        if cpu_base:
            line["cpu_baseline"] = cpu_base
        if parity:
            line["parity"] = parity
        line.update(extra)
        print(json.dumps(line))
    ctx.close()
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()
    return 0


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
    ap.add_argument("--cull", default="off", choices=["on", "off"],
                    help="opt-in per-tile culling of the program (exact: identical frames and counters; reported with executed flops, "
                         "never the headline); paints through the interpreter")
    ap.add_argument("--bind-cpus", default="on", choices=["on", "off"],
                    help="N > 1: bind each rank to the CPUs NVML reports as local to its GPU before it allocates pinned host memory")
    ap.add_argument("--e2e", default="serial", choices=["serial", "pipelined"],
                    help="end-to-end leg of the single-frame workloads.  serial (default): rm_prepare + rm_paint + rm_download per step, "
                         "each step waits for its frame.  pipelined: the steps go through rm_paint_poses_host, which copies frame k to "
                         "host memory while frame k+1 renders (every step still uploads its uniforms and returns its frame)")
    ap.add_argument("--gather", default="p2p", choices=["p2p", "nccl"],
                    help="cfg4, N>1: p2p = every rank stores its strips straight into rank 0's frame over NVLink "
                         "(rm_paint_strips_into); nccl = packed bands + NCCL gather + re-order on rank 0")
    ap.add_argument("--poses", type=int, default=0, help="cfg5: override the 1024-pose batch size")
    ap.add_argument("--strong", default="on", choices=["on", "off"],
                    help="default workload only: also measure the sharded configs cfg4 / cfg5 in this invocation (\"strong\" object)")
    ap.add_argument("--parity", default="on", choices=["on", "off"],
                    help="N=1: compare the rows the cpu_baseline leg renders with the oracle against the timed frame (\"parity\" object)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        return reference_arm(args)
    return b200_arm(args)


if __name__ == "__main__":
    sys.exit(main())
