"""Test-side utilities: ctypes wrapper of the CPU oracle, build-defined scenes S0/S1/S2 and
camera uniforms, written in numpy independently of the product's C++ host so
that the two producers can be checked against each other byte for byte.

Only tests/, bench.py's cpu_baseline leg and __graft_entry__.smoke() may import this module
(it loads oracle/_build/librm_oracle.so).
"""
from __future__ import annotations

import ctypes as C
import math
import os
import subprocess
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
ORACLE_DIR = ROOT / "oracle"
ORACLE_SO = ORACLE_DIR / "_build" / "librm_oracle.so"

# opcodes — /root/reference/src/ray_marching/csg/builder.rs:1-24 (+ extensions X1/X2)
OP_SPHERE, OP_BOX, OP_UNION, OP_SUBTRACTION, OP_INTERSECTION, OP_SMOOTH_UNION = 0, 1, 100, 101, 102, 103
# N4 extension at the values the reference's enum reserves in comments (builder.rs:8,16-23)
OP_PLANE = 2
OP_TRANSLATION_PUSH, OP_TRANSLATION_POP, OP_ROTATION_PUSH, OP_ROTATION_POP, OP_SCALE_PUSH, OP_SCALE_POP = 200, 201, 202, 203, 204, 205
POS_STACK_MAX = 8

DEFAULT_LIMITS = (0.01, 100.0, 100)  # renderer.rs:133-137


def build_oracle(force: bool = False) -> Path:
    src = ORACLE_DIR / "rm_oracle.c"
    if force or not ORACLE_SO.exists() or ORACLE_SO.stat().st_mtime < src.stat().st_mtime:
        subprocess.run(["make", "-C", str(ORACLE_DIR)], check=True, capture_output=True)
    return ORACLE_SO


_lib = None


def oracle_lib():
    global _lib
    if _lib is None:
        build_oracle()
        lib = C.CDLL(str(ORACLE_SO))
        u32p, f32p, u8p, u64p = (C.POINTER(C.c_uint32), C.POINTER(C.c_float), C.POINTER(C.c_uint8),
                                 C.POINTER(C.c_uint64))
        lib.rmo_render.restype = C.c_int
        lib.rmo_render.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, u32p, C.c_uint32, C.c_uint32,
                                   C.c_uint32, C.c_uint32, C.c_uint32, u8p, u8p, f32p, u64p, C.c_int]
        lib.rmo_map.restype = C.c_int
        lib.rmo_map.argtypes = [C.c_void_p, C.c_uint32, u32p, C.c_uint32, f32p, C.c_uint64, f32p]
        lib.rmo_trace.restype = C.c_int
        lib.rmo_trace.argtypes = [C.c_void_p, C.c_uint32, u32p, C.c_uint32, f32p, f32p, f32p]
        lib.rmo_validate.restype = C.c_int
        lib.rmo_validate.argtypes = [C.c_uint32, u32p, C.c_uint32]
        lib.rmo_num_threads.restype = C.c_int
        _lib = lib
    return _lib


def _p(a, ty):
    return a.ctypes.data_as(C.POINTER(ty))


def limits_bytes(min_dist=0.01, max_dist=100.0, max_iter=100) -> bytes:
    return np.array([min_dist, max_dist], np.float32).tobytes() + np.array([max_iter], np.uint32).tobytes()


def _words(words) -> np.ndarray:
    w = np.ascontiguousarray(np.asarray(words, dtype=np.uint32))
    return w if w.size else np.zeros(1, np.uint32)


def oracle_render(uniforms: bytes, cmd_count: int, words, W: int, H: int, row0: int = 0,
                  nrows: int | None = None, limits: bytes | None = None, aov: bool = False,
                  nthreads: int = 0):
    """Returns dict(rgba[nrows,W,4] u8, counters[4] u64, and if aov: cls[nrows,W,16] u8, t[nrows,W,16] f32)."""
    lib = oracle_lib()
    nrows = H - row0 if nrows is None else nrows
    limits = limits_bytes() if limits is None else limits
    n_words = int(np.asarray(words).size)
    w = _words(words)
    rgba = np.zeros((nrows, W, 4), np.uint8)
    cls = np.zeros((nrows, W, 16), np.uint8) if aov else None
    t = np.zeros((nrows, W, 16), np.float32) if aov else None
    cnt = np.zeros(4, np.uint64)
    ub = C.create_string_buffer(uniforms, 144)
    lb = C.create_string_buffer(limits, 12)
    rc = lib.rmo_render(ub, lb, cmd_count, _p(w, C.c_uint32), n_words, W, H, row0, nrows,
                        _p(rgba, C.c_uint8), _p(cls, C.c_uint8) if aov else None,
                        _p(t, C.c_float) if aov else None, _p(cnt, C.c_uint64), nthreads)
    if rc != 0:
        raise RuntimeError(f"rmo_render failed rc={rc}")
    out = {"rgba": rgba, "counters": cnt}
    if aov:
        out["cls"], out["t"] = cls, t
    return out


def oracle_map(cmd_count: int, words, pts: np.ndarray, limits: bytes | None = None) -> np.ndarray:
    lib = oracle_lib()
    limits = limits_bytes() if limits is None else limits
    n_words = int(np.asarray(words).size)
    w = _words(words)
    pts = np.ascontiguousarray(pts, np.float32).reshape(-1, 3)
    out = np.zeros(pts.shape[0], np.float32)
    lb = C.create_string_buffer(limits, 12)
    rc = lib.rmo_map(lb, cmd_count, _p(w, C.c_uint32), n_words, _p(pts, C.c_float), pts.shape[0],
                     _p(out, C.c_float))
    if rc != 0:
        raise RuntimeError(f"rmo_map failed rc={rc}")
    return out


def oracle_trace(cmd_count: int, words, o, d, limits: bytes | None = None):
    lib = oracle_lib()
    limits = limits_bytes() if limits is None else limits
    n_words = int(np.asarray(words).size)
    w = _words(words)
    o = np.asarray(o, np.float32)
    d = np.asarray(d, np.float32)
    out = np.zeros(6, np.float32)
    lb = C.create_string_buffer(limits, 12)
    rc = lib.rmo_trace(lb, cmd_count, _p(w, C.c_uint32), n_words, _p(o, C.c_float), _p(d, C.c_float),
                       _p(out, C.c_float))
    if rc != 0:
        raise RuntimeError(f"rmo_trace failed rc={rc}")
    return {"cls": int(out[0]), "rgb": out[1:4].copy(), "t": float(out[4]), "n_step": int(out[5])}


def oracle_validate(cmd_count: int, words) -> int:
    n_words = int(np.asarray(words).size)
    w = _words(words)
    return oracle_lib().rmo_validate(cmd_count, _p(w, C.c_uint32), n_words)


# ----------------------------------------------------------------------------------------------
# word-buffer construction (wire format of csg/builder.rs:26-62)
# ----------------------------------------------------------------------------------------------
def fbits(x) -> int:
    return int(np.float32(x).view(np.uint32))


def sphere(c, r):
    return [OP_SPHERE, fbits(c[0]), fbits(c[1]), fbits(c[2]), fbits(r)]


def box(c, r):
    return [OP_BOX, fbits(c[0]), fbits(c[1]), fbits(c[2]), fbits(r[0]), fbits(r[1]), fbits(r[2])]


def plane(n, hgt):
    return [OP_PLANE, fbits(n[0]), fbits(n[1]), fbits(n[2]), fbits(hgt)]


def translated(t, child):
    """(words, cmds) of `child` evaluated in a frame moved by t: Push, child, Pop."""
    w, c = child
    return [OP_TRANSLATION_PUSH, fbits(t[0]), fbits(t[1]), fbits(t[2])] + list(w) + [OP_TRANSLATION_POP], c + 2


def rotated(R, child):
    """R: 3x3 rotation matrix (object -> world), sent row-major."""
    w, c = child
    R = np.asarray(R, np.float32).reshape(9)
    return [OP_ROTATION_PUSH] + [fbits(x) for x in R] + list(w) + [OP_ROTATION_POP], c + 2


def scaled(sc, child):
    w, c = child
    return [OP_SCALE_PUSH, fbits(sc)] + list(w) + [OP_SCALE_POP, fbits(sc)], c + 2


def rotation_matrix(axis, angle) -> np.ndarray:
    """Rodrigues formula in float64, rounded to float32 on use."""
    a = np.asarray(axis, np.float64)
    a = a / np.linalg.norm(a)
    K = np.array([[0, -a[2], a[1]], [a[2], 0, -a[0]], [-a[1], a[0], 0]])
    return np.eye(3) + math.sin(angle) * K + (1 - math.cos(angle)) * (K @ K)


def random_program(rng, n_leaves, right_bias=0.3, ext=False, depth=0):
    """Random valid postfix program with mixed tree shapes (value-stack depth <= 32).  ext=True also draws N4
    commands: plane leaves and sub-trees wrapped in translation / rotation / scale groups (nesting <= 8).
    -> (words, cmd_count)."""
    def leaf():
        c = rng.uniform(-2, 2, 3)
        r = rng.random()
        if ext and r < 0.15:
            n = rng.normal(size=3)
            n = (n / np.linalg.norm(n)).astype(np.float32)
            return plane(n, rng.uniform(0.5, 2.5)), 1
        if r < 0.55:
            return sphere(c, rng.uniform(0.2, 1.0)), 1
        return box(c, rng.uniform(0.2, 1.0, 3)), 1

    def op():
        r = rng.random()
        if r < 0.3:
            return [OP_UNION]
        if r < 0.5:
            return [OP_SUBTRACTION]
        if r < 0.65:
            return [OP_INTERSECTION]
        k = [0.25, 0.5, 0.3, 0.1, 1.0][rng.integers(5)]
        return [OP_SMOOTH_UNION, fbits(k)]

    def wrap(node, d):
        r = rng.random()
        if r < 0.4:
            return translated(rng.uniform(-1.5, 1.5, 3), node)
        if r < 0.75:
            return rotated(rotation_matrix(rng.normal(size=3), rng.uniform(-3.1, 3.1)), node)
        return scaled([0.5, 2.0, 0.75, 1.3, 3.0][rng.integers(5)], node)

    def tree(n, d):
        if n == 1:
            node = leaf()
        else:
            nl = 1 if rng.random() > right_bias else int(rng.integers(1, n))
            nl = n - nl if rng.random() < 0.5 else nl
            nl = min(max(nl, 1), n - 1)
            wrap_here = ext and d < POS_STACK_MAX and rng.random() < 0.35
            dd = d + 1 if wrap_here else d
            wa, ca = tree(nl, dd)
            wb, cb = tree(n - nl, dd)
            node = (wa + wb + op(), ca + cb + 1)
            if wrap_here:
                node = wrap(node, d)
            return node
        if ext and d < POS_STACK_MAX and rng.random() < 0.35:
            node = wrap(node, d)
            if d + 1 < POS_STACK_MAX and rng.random() < 0.3:  # a chain of two transforms around one child
                node = wrap(node, d + 1)
        return node

    return tree(n_leaves, depth)


def program_longer_than_its_words(n_leaves: int = 80):
    """A contrived N4 program that decodes to MORE instructions than it has words: a union chain of leaves, each in its
    own Scale group, inside 7 nested Translation groups — every leaf's chain differs from the previous one in its last
    step, so the decoder re-establishes all 8 steps per leaf (11 instructions against 10 words).  -> (words, cmds)"""
    w, c = scaled(1.5, (sphere((0, 0, 0), 0.3), 1))
    for i in range(1, n_leaves):
        lw, lc = scaled(1.5 + 0.01 * i, (sphere((0.1 * i, 0, 0), 0.3), 1))
        w, c = w + lw + [OP_UNION], c + lc + 1
    node = (w, c)
    for d in range(7):
        node = translated((0.01 * d, 0, 0), node)
    return node


def scene_s0():
    """S0: single default sphere (sphere.rs:33-43 template defaults).  -> (cmd_count, words)"""
    return 1, np.array(sphere((0, 0, 0), 1.0), np.uint32)


def scene_s1():
    """S1: Subtraction(Union(Box, Sphere), Sphere) — SURVEY.md §8(d) cfg 2."""
    w = box((0, 0, 0), (1, 1, 1)) + sphere((0, 1, 0), 1.0) + [OP_UNION] + sphere((0.6, 0.3, 1.0), 0.7) + [
        OP_SUBTRACTION]
    return 5, np.array(w, np.uint32)


class SplitMix64:
    def __init__(self, seed: int):
        self.s = seed & 0xFFFFFFFFFFFFFFFF

    def next(self) -> int:
        m = 0xFFFFFFFFFFFFFFFF
        self.s = (self.s + 0x9E3779B97F4A7C15) & m
        z = self.s
        z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & m
        z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & m
        return z ^ (z >> 31)

    def u01(self) -> np.float32:
        return np.float32((self.next() >> 40) * 2.0 ** -24)


def _lerp(lo, hi, u):
    lo, hi = np.float32(lo), np.float32(hi)
    return np.float32(lo + np.float32(np.float32(hi - lo) * u))


def scene_s2(n_leaves: int = 32, seed: int = 0x5EED0040, k: float = 0.25):
    """S2: left-deep chain of 32 leaves, SmoothUnion(k) except Subtraction at i%8==0 — cfg 3/4."""
    rng = SplitMix64(seed)
    w: list[int] = []
    cmds = 0
    for i in range(n_leaves):
        cx = _lerp(-2, 2, rng.u01())
        cy = _lerp(-1, 1.5, rng.u01())
        cz = _lerp(-2, 2, rng.u01())
        if i % 2 == 0:
            r = _lerp(0.25, 0.6, rng.u01())
            w += sphere((cx, cy, cz), r)
        else:
            rx = _lerp(0.2, 0.5, rng.u01())
            ry = _lerp(0.2, 0.5, rng.u01())
            rz = _lerp(0.2, 0.5, rng.u01())
            w += box((cx, cy, cz), (rx, ry, rz))
        cmds += 1
        if i >= 1:
            if i % 8 == 0:
                w += [OP_SUBTRACTION]
            else:
                w += [OP_SMOOTH_UNION, fbits(k)]
            cmds += 1
    return cmds, np.array(w, np.uint32)


# ----------------------------------------------------------------------------------------------
# camera / uniforms (camera.rs:8-13,52-60; renderer.rs:205-222), float64 math rounded to f32
# ----------------------------------------------------------------------------------------------
def orbit_camera(target=(0.0, 0.0, 0.0), radius=5.0, yaw=0.0, pitch=0.0):
    """-> (eye[3], R[3,3]) with R = Ry(-yaw) @ Rx(-pitch), eye = target + R @ (0,0,radius)."""
    a, b = -pitch, -yaw  # roll (about x), pitch (about y) of from_euler_angles(-pitch, -yaw, 0)
    rx = np.array([[1, 0, 0], [0, math.cos(a), -math.sin(a)], [0, math.sin(a), math.cos(a)]])
    ry = np.array([[math.cos(b), 0, math.sin(b)], [0, 1, 0], [-math.sin(b), 0, math.cos(b)]])
    R = ry @ rx
    eye = np.asarray(target, np.float64) + R @ np.array([0.0, 0.0, radius])
    return eye, R


def inv_proj_matrix(aspect: float, fovy=math.pi / 4, near=1.0, far=10000.0) -> np.ndarray:
    """Inverse of nalgebra Perspective3::new(aspect, fovy, near, far) (renderer.rs:206-210)."""
    t = math.tan(fovy / 2)
    m11, m22 = 1.0 / (aspect * t), 1.0 / t
    m33, m34 = (far + near) / (near - far), 2 * far * near / (near - far)
    P = np.array([[m11, 0, 0, 0], [0, m22, 0, 0], [0, 0, m33, m34], [0, 0, -1, 0]], np.float64)
    return np.linalg.inv(P)


def uniforms_bytes(W: float, H: float, eye, R) -> bytes:
    """144-byte Uniforms block: extent @0, inv_proj (column-major) @16, inv_view @80."""
    inv_view = np.eye(4)
    inv_view[:3, :3] = R
    inv_view[:3, 3] = eye
    ip = inv_proj_matrix(W / H)
    buf = np.zeros(36, np.float32)
    buf[0], buf[1] = W, H
    buf[4:20] = ip.T.reshape(-1).astype(np.float32)  # column-major = transpose flattened
    buf[20:36] = inv_view.T.reshape(-1).astype(np.float32)
    return buf.tobytes()


def orbit_uniforms(W: int, H: int, yaw=0.0, pitch=0.0, radius=5.0, target=(0, 0, 0)) -> bytes:
    eye, R = orbit_camera(target, radius, yaw, pitch)
    return uniforms_bytes(float(W), float(H), eye, R)


def scene_s3(n_inst: int = 8, k: float = 0.25):
    """S3 (N4): ring of rotated + translated instances, every other one scaled, over the plane y = -1
    (definition: ray-marching_b200/scenes.py node_s3; built here independently from the word builders)."""
    w: list[int] = []
    cmds = 0
    for i in range(n_inst):
        if i % 2 == 0:
            leaf = (box((0, 0, 0), (0.35, 0.5, 0.2)), 1)
        else:
            leaf = scaled(0.8, (sphere((0, 0, 0), 0.6) + box((0, 0, 0), (0.7, 0.2, 0.2)) + [OP_SUBTRACTION], 3))
        a = 2 * math.pi * i / n_inst
        R = np.array([[math.cos(a), 0, math.sin(a)], [0, 1, 0], [-math.sin(a), 0, math.cos(a)]])
        ty = np.float32(0.1) * np.float32(i) - np.float32(0.3)
        iw, ic = rotated(R, translated((1.6, ty, 0.0), leaf))
        w += iw
        cmds += ic
        if i >= 1:
            w += [OP_SMOOTH_UNION, fbits(k)]
            cmds += 1
    w += plane((0, 1, 0), 1.0) + [OP_UNION]
    return cmds + 2, np.array(w, np.uint32)


SCENES = {"S0": scene_s0, "S1": scene_s1, "S2": scene_s2, "S3": scene_s3}
