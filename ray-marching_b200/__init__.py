"""ray_marching_b200 — ctypes binding of the C-ABI render-pass library (include/rm_b200.h).

The directory is named ``ray-marching_b200``; import it with ``rm_import.load()`` (repo root) which
registers it as the module ``ray_marching_b200``.

There is exactly one implementation behind this binding: the sm_100a CUDA kernels in
``librm_b200.so``.  If the library is missing, or there is no CUDA device, calls raise — there is no
CPU path (the CPU oracle lives in ``oracle/`` and is test infrastructure only).
"""
from __future__ import annotations

import ctypes as C
import subprocess
from pathlib import Path

import numpy as np

PKG_DIR = Path(__file__).resolve().parent
LIB_PATH = PKG_DIR / "librm_b200.so"
HOST_LIB_PATH = PKG_DIR / "librm_host.so"

RM_OK = 0
E_PROGRAM_TOO_LARGE, E_BAD_OPCODE, E_STACK_OVERFLOW, E_STACK_UNBALANCED, E_BAD_PARAM, E_JIT, E_UNSUPPORTED = -2, -3, -4, -5, -8, -9, -10
# command opcodes (include/rm_b200.h): the reference's four, the two extension operators, and the N4 commands
OP_SPHERE, OP_BOX, OP_PLANE, OP_UNION, OP_SUBTRACTION, OP_INTERSECTION, OP_SMOOTH_UNION = 0, 1, 2, 100, 101, 102, 103
OP_TRANSLATION_PUSH, OP_TRANSLATION_POP, OP_ROTATION_PUSH, OP_ROTATION_POP, OP_SCALE_PUSH, OP_SCALE_POP = 200, 201, 202, 203, 204, 205
MODE_EXACT, MODE_FAST = 0, 1
SPEC_LITERAL_TIER = 0x100  # or-ed into the mode of spec_source / spec_compile_check: the literal tier's kernel
KERNEL_WAVEFRONT, KERNEL_SIMPLE, KERNEL_WAVEFRONT2, KERNEL_DIRECT = 0, 1, 2, 3
UNIFORMS_BYTES, LIMITS_BYTES = 144, 12

# every symbol include/rm_b200.h declares (checked by tests/test_abi_cpu.py)
ABI_SYMBOLS = [
    "rm_create", "rm_create_ex", "rm_destroy", "rm_set_limits", "rm_set_mode", "rm_get_mode", "rm_set_kernel", "rm_set_early_out",
    "rm_prepare", "rm_prepare_uniforms", "rm_program_uploads", "rm_validate_program", "rm_paint", "rm_paint_strips", "rm_paint_strips_into", "rm_shared_frame_create", "rm_shared_frame_open",
    "rm_shared_frame_close", "rm_copy_to_host", "rm_copy_strips_to_host", "rm_host_register", "rm_host_unregister", "rm_paint_poses", "rm_paint_poses_host", "rm_download",
    "rm_sync", "rm_paint_debug", "rm_paint_aov", "rm_map", "rm_set_counters", "rm_get_counters", "rm_mark", "rm_elapsed_ms",
    "rm_flush_l2", "rm_fp32_peak", "rm_launch_count", "rm_host_alloc", "rm_host_free", "rm_last_error",
    "rm_strerror", "rm_version",
    "rm_group_create", "rm_group_destroy", "rm_group_size", "rm_group_ctx", "rm_group_set_mode", "rm_group_set_limits",
    "rm_specialize", "rm_set_specialize", "rm_spec_info", "rm_spec_status", "rm_spec_source", "rm_spec_compile_check", "rm_spec_params", "rm_spec_cache", "rm_set_spec_literal_after", "rm_spec_tier", "rm_set_cull", "rm_get_cull_stats", "rm_paint_culled_static",
    "rm_group_specialize", "rm_group_prepare", "rm_group_paint_bands", "rm_group_frame", "rm_group_paint_poses", "rm_group_last_error",
]


class RmError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"rm_b200 error {code}: {msg}")
        self.code = code


def build(verbose: bool = False) -> None:
    """Compile librm_b200.so and librm_host.so in-tree for sm_100a (nvcc cross-compiles without a GPU)."""
    for d in ("csrc", "host"):
        if not (PKG_DIR / d / "Makefile").exists():
            continue
        r = subprocess.run(["make", "-C", str(PKG_DIR / d)], capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"build of {d} failed:\n{r.stdout}\n{r.stderr}")
        if verbose:
            print(r.stdout)


_lib = None


def lib() -> C.CDLL:
    """Loads librm_b200.so; raises if it has not been built (no fallback)."""
    global _lib
    if _lib is None:
        if not LIB_PATH.exists():
            raise RuntimeError(f"{LIB_PATH} is missing: run __graft_entry__.build() (nvcc, sm_100a) first")
        L = C.CDLL(str(LIB_PATH))
        vp, u32, u64, sz, i = C.c_void_p, C.c_uint32, C.c_uint64, C.c_size_t, C.c_int
        sig = {
            "rm_create": (i, [i, C.POINTER(vp)]),
            "rm_create_ex": (i, [i, u32, C.POINTER(vp)]),
            "rm_destroy": (None, [vp]),
            "rm_set_limits": (i, [vp, vp]),
            "rm_set_early_out": (i, [vp, i]),
            "rm_set_mode": (i, [vp, i]),
            "rm_get_mode": (i, [vp]),
            "rm_set_kernel": (i, [vp, i]),
            "rm_prepare": (i, [vp, vp, u32, vp, u32]),
            "rm_prepare_uniforms": (i, [vp, vp]),
            "rm_validate_program": (i, [u32, vp, u32, u32]),
            "rm_paint": (i, [vp, u32, u32, u32, u32, vp, sz, vp]),
            "rm_paint_poses": (i, [vp, vp, u32, u32, u32, vp, sz, sz, vp]),
            "rm_paint_poses_host": (i, [vp, vp, u32, u32, u32, vp, sz, u32]),
            "rm_program_uploads": (u64, [vp]),
            "rm_paint_strips": (i, [vp, u32, u32, u32, u32, u32, vp, sz, vp, C.POINTER(u32)]),
            "rm_paint_strips_into": (i, [vp, u32, u32, u32, u32, u32, vp, sz, vp, C.POINTER(u32)]),
            "rm_shared_frame_create": (i, [vp, sz, C.POINTER(vp), vp]),
            "rm_shared_frame_open": (i, [vp, vp, C.POINTER(vp)]),
            "rm_shared_frame_close": (i, [vp, vp]),
            "rm_copy_to_host": (i, [vp, vp, vp, sz, vp]),
            "rm_copy_strips_to_host": (i, [vp, vp, u32, u32, u32, u32, u32, vp, vp]),
            "rm_host_register": (i, [vp, sz]),
            "rm_host_unregister": (i, [vp]),
            "rm_download": (i, [vp, vp, sz]),
            "rm_sync": (i, [vp]),
            "rm_paint_debug": (i, [vp, u32, u32, u32, u32, vp, vp, vp, vp]),
            "rm_paint_aov": (i, [vp, u32, u32, u32, u32, vp, vp, vp, C.POINTER(i)]),
            "rm_map": (i, [vp, vp, u64, vp]),
            "rm_set_counters": (i, [vp, i]),
            "rm_get_counters": (i, [vp, vp, i]),
            "rm_mark": (i, [vp, i]),
            "rm_elapsed_ms": (i, [vp, i, i, C.POINTER(C.c_float)]),
            "rm_flush_l2": (i, [vp]),
            "rm_fp32_peak": (i, [vp, C.POINTER(C.c_double), C.POINTER(i), C.POINTER(i)]),
            "rm_launch_count": (u64, [vp]),
            "rm_host_alloc": (vp, [sz]),
            "rm_host_free": (None, [vp]),
            "rm_last_error": (C.c_char_p, [vp]),
            "rm_strerror": (C.c_char_p, [i]),
            "rm_version": (C.c_char_p, []),
            "rm_specialize": (i, [vp]),
            "rm_set_specialize": (i, [vp, i]),
            "rm_spec_info": (i, [vp, C.POINTER(i), C.POINTER(i), C.POINTER(i), C.POINTER(C.c_double), C.POINTER(u64),
                                 C.POINTER(u64), C.POINTER(u64)]),
            "rm_spec_status": (C.c_char_p, [vp]),
            "rm_spec_source": (i, [u32, vp, u32, i, i, C.c_char_p, sz, C.POINTER(sz)]),
            "rm_spec_compile_check": (i, [u32, vp, u32, i, i, C.POINTER(u64), C.POINTER(C.c_double), C.c_char_p, sz]),
            "rm_spec_params": (i, [u32, vp, u32, i, vp, sz, C.POINTER(sz)]),
            "rm_set_cull": (i, [vp, i]),
            "rm_paint_culled_static": (i, [vp, u32, u32, u32, u32, vp, sz, vp]),
            "rm_get_cull_stats": (i, [vp, C.POINTER(u64), C.POINTER(u64), i]),
            "rm_set_spec_literal_after": (i, [vp, u32]),
            "rm_spec_tier": (i, [vp]),
            "rm_spec_cache": (i, [u64, C.POINTER(u64), C.POINTER(u64), C.POINTER(u64)]),
            "rm_group_specialize": (i, [vp]),
            "rm_group_create": (i, [C.POINTER(i), i, C.POINTER(vp)]),
            "rm_group_destroy": (None, [vp]),
            "rm_group_size": (i, [vp]),
            "rm_group_ctx": (vp, [vp, i]),
            "rm_group_set_mode": (i, [vp, i]),
            "rm_group_set_limits": (i, [vp, vp]),
            "rm_group_prepare": (i, [vp, vp, u32, vp, u32]),
            "rm_group_paint_bands": (i, [vp, u32, u32, u32, vp]),
            "rm_group_frame": (i, [vp, C.POINTER(vp), C.POINTER(sz)]),
            "rm_group_paint_poses": (i, [vp, vp, u32, u32, u32, vp, sz]),
            "rm_group_last_error": (C.c_char_p, [vp]),
        }
        for name, (res, args) in sig.items():
            f = getattr(L, name)
            f.restype, f.argtypes = res, args
        _lib = L
    return _lib


def _words_arr(words) -> np.ndarray:
    return np.ascontiguousarray(np.asarray(words, dtype=np.uint32))


def validate_program(cmd_count: int, words, max_words: int = 255) -> int:
    """Static validation only — no device needed.  Returns an RM_* code."""
    w = _words_arr(words)
    return lib().rm_validate_program(cmd_count, w.ctypes.data if w.size else None, w.size, max_words)


def spec_source(cmd_count: int, words, mode: int = MODE_FAST, rays_per_thread: int = 0) -> str:
    """CUDA source of the specialised kernel for a command stream (no device needed).  Raises RmError."""
    w = _words_arr(words)
    need = C.c_size_t()
    rc = lib().rm_spec_source(cmd_count, w.ctypes.data if w.size else None, w.size, mode, rays_per_thread, None, 0, C.byref(need))
    if rc != RM_OK:
        raise RmError(rc, lib().rm_strerror(rc).decode())
    buf = C.create_string_buffer(need.value)
    lib().rm_spec_source(cmd_count, w.ctypes.data if w.size else None, w.size, mode, rays_per_thread, buf, need.value, None)
    return buf.value.decode()


def spec_params(cmd_count: int, words, mode: int = MODE_FAST) -> np.ndarray:
    """The parameter block (float32) the generated kernel of these words reads: its non-zero parameters in stream order."""
    w = _words_arr(words)
    need = C.c_size_t()
    rc = lib().rm_spec_params(cmd_count, w.ctypes.data if w.size else None, w.size, mode, None, 0, C.byref(need))
    if rc != RM_OK:
        raise RmError(rc, lib().rm_strerror(rc).decode())
    out = np.zeros(max(1, need.value), np.float32)
    lib().rm_spec_params(cmd_count, w.ctypes.data if w.size else None, w.size, mode, out.ctypes.data, out.size, None)
    return out[:need.value]


def spec_cache(new_capacity: int = 0) -> dict:
    """Process-wide specialised-kernel cache: optionally sets the LRU bound; -> {entries, capacity, evictions}."""
    n, cap, ev = C.c_uint64(), C.c_uint64(), C.c_uint64()
    lib().rm_spec_cache(new_capacity, C.byref(n), C.byref(cap), C.byref(ev))
    return {"entries": int(n.value), "capacity": int(cap.value), "evictions": int(ev.value)}


def spec_compile_check(cmd_count: int, words, mode: int = MODE_FAST, rays_per_thread: int = 0) -> dict:
    """Trial NVRTC compile for sm_100a (no device needed) -> {rc, cubin_bytes, compile_ms, log}."""
    w = _words_arr(words)
    cb, ms, log = C.c_uint64(), C.c_double(), C.create_string_buffer(16384)
    rc = lib().rm_spec_compile_check(cmd_count, w.ctypes.data if w.size else None, w.size, mode, rays_per_thread,
                                     C.byref(cb), C.byref(ms), log, 16384)
    return {"rc": rc, "cubin_bytes": int(cb.value), "compile_ms": ms.value, "log": log.value.decode(errors="replace")}


def host_register(ptr: int, nbytes: int) -> None:
    """Page-locks caller-owned host memory (e.g. a shared-memory mapping) so that device copies into it are DMA."""
    rc = lib().rm_host_register(ptr, nbytes)
    if rc != RM_OK:
        raise RmError(rc, "rm_host_register failed")


def host_unregister(ptr: int) -> None:
    rc = lib().rm_host_unregister(ptr)
    if rc != RM_OK:
        raise RmError(rc, "rm_host_unregister failed")


class PinnedBuffer:
    """Page-locked host buffer exposed as a numpy uint8 array."""

    def __init__(self, nbytes: int):
        self.ptr = lib().rm_host_alloc(nbytes)
        if not self.ptr:
            raise RmError(-6, "rm_host_alloc failed")
        self.nbytes = nbytes
        self.array = np.ctypeslib.as_array((C.c_uint8 * nbytes).from_address(self.ptr))

    def free(self):
        if self.ptr:
            lib().rm_host_free(self.ptr)
            self.ptr = None

    def __del__(self):
        self.free()


class Context:
    """One device's render-pass resources (the reference's RayMarchingResources, renderer.rs:43-175)."""

    def __init__(self, device: int = 0, max_words: int = 255):
        self._L = lib()
        h = C.c_void_p()
        rc = self._L.rm_create_ex(device, max_words, C.byref(h))
        if rc != RM_OK:
            raise RmError(rc, f"rm_create_ex(device={device}): {self._L.rm_strerror(rc).decode()} "
                              "(a CUDA device of compute capability 10.x is required; there is no CPU path)")
        self._h = h
        self.device = device

    def close(self):
        if getattr(self, "_h", None):
            self._L.rm_destroy(self._h)
            self._h = None

    def __del__(self):
        self.close()

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def _chk(self, rc: int):
        if rc != RM_OK:
            raise RmError(rc, self._L.rm_last_error(self._h).decode() or self._L.rm_strerror(rc).decode())

    # ---- configuration
    def set_limits(self, min_dist=0.01, max_dist=100.0, max_iter=100):
        b = np.array([min_dist, max_dist], np.float32).tobytes() + np.array([max_iter], np.uint32).tobytes()
        self._chk(self._L.rm_set_limits(self._h, b))

    def set_mode(self, mode: int):
        self._chk(self._L.rm_set_mode(self._h, mode))

    def set_kernel(self, kernel: int):
        self._chk(self._L.rm_set_kernel(self._h, kernel))

    # ---- prepare (renderer.rs:196-242)
    def set_early_out(self, on: bool):
        self._chk(self._L.rm_set_early_out(self._h, int(bool(on))))

    def set_cull(self, on: bool):
        """Opt-in per-tile culling of the program (exact: identical frames; reported separately, never the headline)."""
        self._chk(self._L.rm_set_cull(self._h, int(bool(on))))

    def paint_culled_static(self, W: int, H: int, row0: int, nrows: int, dst_ptr: int, pitch: int = 0, stream: int | None = None):
        """Allocation-free culled paint into the caller's device buffer (masks in a static device array; rm_b200.h)."""
        self._chk(self._L.rm_paint_culled_static(self._h, W, H, row0, nrows, dst_ptr, pitch or W * 4, stream))

    def cull_stats(self, reset: bool = True) -> dict:
        f, n = C.c_uint64(), C.c_uint64()
        self._chk(self._L.rm_get_cull_stats(self._h, C.byref(f), C.byref(n), int(reset)))
        return {"executed_map_flops": int(f.value), "culled_paints": int(n.value)}

    # ---- program specialisation (run-time compiled straight-line map_scene)
    def specialize(self):
        """Compile the prepared program for the current mode now; raises RmError(-9) if that is not possible."""
        self._chk(self._L.rm_specialize(self._h))

    def set_specialize(self, on):
        """0 / False: interpreter; 1 / True: compile every program at its first paint; 2: compile in the background."""
        self._chk(self._L.rm_set_specialize(self._h, int(on)))

    def set_spec_literal_after(self, n_paints: int):
        """Literal tier: adopted after n_paints plain paints of unchanged words (0 = structure tier only; default 8)."""
        self._chk(self._L.rm_set_spec_literal_after(self._h, n_paints))

    @property
    def spec_tier(self) -> int:
        """What the next plain paint runs: 0 interpreter, 1 structure-specialised kernel, 2 literal-specialised kernel."""
        return int(self._L.rm_spec_tier(self._h))

    def spec_info(self) -> dict:
        a, r, mb, ms = C.c_int(), C.c_int(), C.c_int(), C.c_double()
        cb, nc, nl = C.c_uint64(), C.c_uint64(), C.c_uint64()
        self._chk(self._L.rm_spec_info(self._h, C.byref(a), C.byref(r), C.byref(mb), C.byref(ms), C.byref(cb), C.byref(nc), C.byref(nl)))
        return {"active": bool(a.value), "rays_per_thread": r.value, "min_blocks": mb.value, "compile_ms": ms.value,
                "cubin_bytes": int(cb.value), "compiles": int(nc.value), "launches": int(nl.value),
                "status": self._L.rm_spec_status(self._h).decode(errors="replace")}

    def prepare(self, uniforms: bytes, cmd_count: int, words):
        assert len(uniforms) == UNIFORMS_BYTES
        w = _words_arr(words)
        self._chk(self._L.rm_prepare(self._h, uniforms, cmd_count, w.ctypes.data if w.size else None, w.size))

    @property
    def program_uploads(self) -> int:
        return int(self._L.rm_program_uploads(self._h))

    def prepare_uniforms(self, uniforms: bytes):
        assert len(uniforms) == UNIFORMS_BYTES
        self._chk(self._L.rm_prepare_uniforms(self._h, uniforms))

    # ---- paint (renderer.rs:244-255)
    def paint(self, W: int, H: int, row0: int = 0, nrows: int | None = None, dst_ptr: int | None = None,
              pitch: int = 0, stream: int | None = None):
        nrows = H - row0 if nrows is None else nrows
        self._chk(self._L.rm_paint(self._h, W, H, row0, nrows, dst_ptr, pitch, stream))

    def paint_strips(self, W: int, H: int, strip_rows: int, first_strip: int, strip_step: int,
                     dst_ptr: int | None = None, pitch: int = 0, stream: int | None = None) -> int:
        """Shades strips first_strip, first_strip+strip_step, ... packed contiguously; returns rows painted."""
        n = C.c_uint32()
        self._chk(self._L.rm_paint_strips(self._h, W, H, strip_rows, first_strip, strip_step, dst_ptr, pitch, stream,
                                          C.byref(n)))
        return int(n.value)

    def paint_strips_into(self, W: int, H: int, strip_rows: int, first_strip: int, strip_step: int, frame_ptr: int,
                          pitch: int | None = None, stream: int | None = None) -> int:
        """This shard's strips stored at their frame rows in a whole-frame buffer (possibly another GPU's)."""
        n = C.c_uint32()
        self._chk(self._L.rm_paint_strips_into(self._h, W, H, strip_rows, first_strip, strip_step, frame_ptr,
                                               pitch or W * 4, stream, C.byref(n)))
        return int(n.value)

    def shared_frame_create(self, nbytes: int) -> tuple[int, bytes]:
        """-> (device pointer, 64-byte handle to send to the other processes)."""
        p, hnd = C.c_void_p(), C.create_string_buffer(64)
        self._chk(self._L.rm_shared_frame_create(self._h, nbytes, C.byref(p), hnd))
        return int(p.value), hnd.raw

    def shared_frame_open(self, handle: bytes) -> int:
        assert len(handle) == 64
        p = C.c_void_p()
        self._chk(self._L.rm_shared_frame_open(self._h, C.create_string_buffer(handle, 64), C.byref(p)))
        return int(p.value)

    def shared_frame_close(self, ptr: int):
        self._chk(self._L.rm_shared_frame_close(self._h, ptr))

    def copy_to_host(self, dev_ptr: int, out, nbytes: int | None = None, stream: int | None = None):
        """Device -> host copy of any addressable device pointer into a PinnedBuffer or a C-contiguous array."""
        if isinstance(out, PinnedBuffer):
            self._chk(self._L.rm_copy_to_host(self._h, dev_ptr, out.ptr, nbytes or out.nbytes, stream))
            return out.array
        assert out.flags["C_CONTIGUOUS"]
        self._chk(self._L.rm_copy_to_host(self._h, dev_ptr, out.ctypes.data, nbytes or out.nbytes, stream))
        return out

    def copy_strips_to_host(self, band_ptr: int | None, W: int, H: int, strip_rows: int, first_strip: int, strip_step: int,
                            host_frame_ptr: int, stream: int | None = None):
        """Packed strips of this shard (band_ptr None = the ctx-owned buffer) -> their rows of a whole host frame."""
        self._chk(self._L.rm_copy_strips_to_host(self._h, band_ptr, W, H, strip_rows, first_strip, strip_step, host_frame_ptr, stream))

    def paint_poses(self, uniforms_array: bytes, n_poses: int, W: int, H: int, dst_ptr: int | None = None,
                    pitch: int = 0, frame_stride: int = 0, stream: int | None = None):
        assert len(uniforms_array) >= n_poses * UNIFORMS_BYTES
        self._chk(self._L.rm_paint_poses(self._h, uniforms_array, n_poses, W, H, dst_ptr, pitch, frame_stride, stream))

    def paint_poses_host(self, uniforms_array: bytes, n_poses: int, W: int, H: int, host_ptr: int,
                         host_frame_stride: int | None = None, host_ring_frames: int = 0):
        """Pose batch with every frame returned to host memory (D2H of frame p overlaps the render of p+1)."""
        assert len(uniforms_array) >= n_poses * UNIFORMS_BYTES
        self._chk(self._L.rm_paint_poses_host(self._h, uniforms_array, n_poses, W, H, host_ptr,
                                              host_frame_stride or W * H * 4, host_ring_frames))

    def download(self, out: np.ndarray | PinnedBuffer | None = None, nbytes: int | None = None) -> np.ndarray:
        if isinstance(out, PinnedBuffer):
            self._chk(self._L.rm_download(self._h, out.ptr, nbytes or out.nbytes))
            return out.array
        assert out is not None and out.flags["C_CONTIGUOUS"]
        self._chk(self._L.rm_download(self._h, out.ctypes.data, nbytes or out.nbytes))
        return out

    def render(self, uniforms: bytes, cmd_count: int, words, W: int, H: int, row0: int = 0,
               nrows: int | None = None) -> np.ndarray:
        """prepare + paint + download -> uint8 [nrows, W, 4] (host)."""
        nrows = H - row0 if nrows is None else nrows
        self.prepare(uniforms, cmd_count, words)
        self.paint(W, H, row0, nrows)
        out = np.empty((nrows, W, 4), np.uint8)
        if nrows:
            self.download(out)
        return out

    def sync(self):
        self._chk(self._L.rm_sync(self._h))

    # ---- verification surface
    def paint_debug(self, W: int, H: int, row0: int = 0, nrows: int | None = None, aov: bool = True):
        nrows = H - row0 if nrows is None else nrows
        cls = np.zeros((nrows, W, 16), np.uint8) if aov else None
        t = np.zeros((nrows, W, 16), np.float32) if aov else None
        cnt = np.zeros(4, np.uint64)
        rgba = np.zeros((nrows, W, 4), np.uint8)
        self._chk(self._L.rm_paint_debug(self._h, W, H, row0, nrows, cls.ctypes.data if aov else None,
                                         t.ctypes.data if aov else None, cnt.ctypes.data, rgba.ctypes.data))
        out = {"rgba": rgba, "counters": cnt}
        if aov:
            out["cls"], out["t"] = cls, t
        return out

    def paint_aov(self, W: int, H: int, row0: int = 0, nrows: int | None = None) -> dict:
        """Image + per-sample class / distance from the production kernel itself (specialised if that is on)."""
        nrows = H - row0 if nrows is None else nrows
        cls = np.zeros((nrows, W, 16), np.uint8)
        t = np.zeros((nrows, W, 16), np.float32)
        rgba = np.zeros((nrows, W, 4), np.uint8)
        sp = C.c_int()
        self._chk(self._L.rm_paint_aov(self._h, W, H, row0, nrows, cls.ctypes.data, t.ctypes.data, rgba.ctypes.data, C.byref(sp)))
        return {"rgba": rgba, "cls": cls, "t": t, "specialised": bool(sp.value)}

    def map(self, pts: np.ndarray) -> np.ndarray:
        pts = np.ascontiguousarray(pts, np.float32).reshape(-1, 3)
        out = np.zeros(pts.shape[0], np.float32)
        self._chk(self._L.rm_map(self._h, pts.ctypes.data, pts.shape[0], out.ctypes.data))
        return out

    def set_counters(self, on: bool):
        self._chk(self._L.rm_set_counters(self._h, int(on)))

    def get_counters(self, reset: bool = True) -> np.ndarray:
        cnt = np.zeros(4, np.uint64)
        self._chk(self._L.rm_get_counters(self._h, cnt.ctypes.data, int(reset)))
        return cnt

    # ---- measurement
    def mark(self, slot: int):
        self._chk(self._L.rm_mark(self._h, slot))

    def elapsed_ms(self, a: int, b: int) -> float:
        ms = C.c_float()
        self._chk(self._L.rm_elapsed_ms(self._h, a, b, C.byref(ms)))
        return ms.value

    def flush_l2(self):
        self._chk(self._L.rm_flush_l2(self._h))

    def fp32_peak(self):
        tf, sm, khz = C.c_double(), C.c_int(), C.c_int()
        self._chk(self._L.rm_fp32_peak(self._h, C.byref(tf), C.byref(sm), C.byref(khz)))
        return {"tflops": tf.value, "sm_count": sm.value, "sm_clock_khz": khz.value}

    @property
    def launch_count(self) -> int:
        return int(self._L.rm_launch_count(self._h))


class Group:
    """One host thread driving several GPUs (include/rm_b200.h rm_group_*): one Context per device, the frame of
    paint_bands lives on the first device and the others store into it over NVLink."""

    def __init__(self, devices):
        self._L = lib()
        self.devices = [int(d) for d in devices]
        arr = (C.c_int * len(self.devices))(*self.devices)
        h = C.c_void_p()
        rc = self._L.rm_group_create(arr, len(self.devices), C.byref(h))
        if rc != RM_OK:
            raise RmError(rc, f"rm_group_create({self.devices}): {self._L.rm_strerror(rc).decode()} "
                              "(every device must be sm_100 and able to address the first one)")
        self._h = h

    def close(self):
        if getattr(self, "_h", None):
            self._L.rm_group_destroy(self._h)
            self._h = None

    def __del__(self):
        self.close()

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def __len__(self):
        return int(self._L.rm_group_size(self._h))

    def _chk(self, rc: int):
        if rc != RM_OK:
            raise RmError(rc, self._L.rm_group_last_error(self._h).decode() or self._L.rm_strerror(rc).decode())

    def member_launch_count(self, i: int) -> int:
        return int(self._L.rm_launch_count(self._L.rm_group_ctx(self._h, i)))

    def set_kernel(self, kernel: int):
        for i in range(len(self)):
            rc = self._L.rm_set_kernel(self._L.rm_group_ctx(self._h, i), kernel)
            if rc != RM_OK:
                raise RmError(rc, self._L.rm_strerror(rc).decode())

    def set_mode(self, mode: int):
        self._chk(self._L.rm_group_set_mode(self._h, mode))

    def set_limits(self, min_dist=0.01, max_dist=100.0, max_iter=100):
        b = np.array([min_dist, max_dist], np.float32).tobytes() + np.array([max_iter], np.uint32).tobytes()
        self._chk(self._L.rm_group_set_limits(self._h, b))

    def prepare(self, uniforms: bytes, cmd_count: int, words):
        assert len(uniforms) == UNIFORMS_BYTES
        w = _words_arr(words)
        self._chk(self._L.rm_group_prepare(self._h, uniforms, cmd_count, w.ctypes.data if w.size else None, w.size))

    def specialize(self):
        self._chk(self._L.rm_group_specialize(self._h))

    def paint_bands(self, W: int, H: int, strip_rows: int = 16, out=None):
        """ONE frame across the members; `out` (PinnedBuffer or C-contiguous array of W*H*4 bytes) receives it."""
        ptr = None
        if isinstance(out, PinnedBuffer):
            ptr = out.ptr
        elif out is not None:
            assert out.flags["C_CONTIGUOUS"] and out.nbytes >= W * H * 4
            ptr = out.ctypes.data
        self._chk(self._L.rm_group_paint_bands(self._h, W, H, strip_rows, ptr))
        return out.array if isinstance(out, PinnedBuffer) else out

    def frame(self) -> tuple[int, int]:
        p, pitch = C.c_void_p(), C.c_size_t()
        self._chk(self._L.rm_group_frame(self._h, C.byref(p), C.byref(pitch)))
        return int(p.value), int(pitch.value)

    def paint_poses(self, uniforms_array: bytes, n_poses: int, W: int, H: int, host_ptr: int,
                    host_frame_stride: int | None = None):
        assert len(uniforms_array) >= n_poses * UNIFORMS_BYTES
        self._chk(self._L.rm_group_paint_poses(self._h, uniforms_array, n_poses, W, H, host_ptr,
                                               host_frame_stride or W * H * 4))
