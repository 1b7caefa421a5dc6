"""Build-defined benchmark scenes S0/S1/S2 constructed through the product's C++
host (CSGNode constructors + BuildCommands flattener), and the orbit-camera uniforms of each config.
tests/helpers.py builds the same scenes independently in numpy; the two are compared byte for byte."""
from __future__ import annotations

import math

import numpy as np

from . import host


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


def node_s0() -> int:
    """single sphere with the template defaults (sphere.rs:33-43)"""
    return host.sphere((0, 0, 0), 1.0)


def node_s1() -> int:
    """Subtraction(Union(Box, Sphere), Sphere)"""
    return host.scene_s1()


def node_s2(n_leaves: int = 32, seed: int = 0x5EED0040, k: float = 0.25) -> int:
    """left-deep chain: SmoothUnion(k) except Subtraction at i % 8 == 0; spheres at even i, boxes at odd i"""
    rng = SplitMix64(seed)
    acc = None
    for i in range(n_leaves):
        cx, cy, cz = _lerp(-2, 2, rng.u01()), _lerp(-1, 1.5, rng.u01()), _lerp(-2, 2, rng.u01())
        if i % 2 == 0:
            leaf = host.sphere((cx, cy, cz), _lerp(0.25, 0.6, rng.u01()))
        else:
            leaf = host.box((cx, cy, cz), [_lerp(0.2, 0.5, rng.u01()) for _ in range(3)])
        if acc is None:
            acc = leaf
        elif i % 8 == 0:
            acc = host.binary(host.SUBTRACTION, acc, leaf)
        else:
            acc = host.binary(host.SMOOTH_UNION, acc, leaf, k)
    return acc


def _rot_y(angle: float) -> np.ndarray:
    """row-major rotation about +y (object -> world), float64 rounded to float32"""
    c, s = math.cos(angle), math.sin(angle)
    return np.array([[c, 0, s], [0, 1, 0], [-s, 0, c]], np.float64).astype(np.float32)


def node_s3(n_inst: int = 8, k: float = 0.25) -> int:
    """N4 scene (plane + space transforms): a ring of n_inst instances — instance i is Rotation(y, 2*pi*i/n) of
    Translation((1.6, 0.1*i - 0.3, 0)) of a leaf in its own frame (even i: Box r=(0.35,0.5,0.2); odd i: Scale(0.8) of
    Subtraction(Sphere r=0.6, Box r=(0.7,0.2,0.2)) — smooth-united left to right, then Union with the plane y = -1."""
    acc = None
    for i in range(n_inst):
        if i % 2 == 0:
            leaf = host.box((0, 0, 0), (0.35, 0.5, 0.2))
        else:
            leaf = host.scale(host.binary(host.SUBTRACTION, host.sphere((0, 0, 0), 0.6), host.box((0, 0, 0), (0.7, 0.2, 0.2))), 0.8)
        inst = host.rotation_matrix(host.translation(leaf, (1.6, np.float32(0.1) * np.float32(i) - np.float32(0.3), 0.0)),
                                    _rot_y(2 * math.pi * i / n_inst))
        acc = inst if acc is None else host.binary(host.SMOOTH_UNION, acc, inst, k)
    return host.binary(host.UNION, acc, host.plane((0, 1, 0), 1.0))


NODES = {"S0": node_s0, "S1": node_s1, "S2": node_s2, "S3": node_s3}


def program(scene: str):
    """-> (cmd_count, words) of a scene, flattened by the C++ host."""
    return host.build_commands(NODES[scene]())


def orbit_uniforms(W: int, H: int, yaw: float = 0.0, pitch: float = 0.0, radius: float = 5.0, target=(0, 0, 0)) -> bytes:
    o = host.Orbit(target, radius)
    o.set(yaw, pitch)
    return o.uniforms(W, H)


def orbit_pose_batch(W: int, H: int, n: int, pitch: float = 0.4, radius: float = 5.0) -> bytes:
    """config 5: yaw_k = 2*pi*k/n, k = 0..n-1"""
    return b"".join(orbit_uniforms(W, H, yaw=2 * math.pi * k / n, pitch=pitch, radius=radius) for k in range(n))
