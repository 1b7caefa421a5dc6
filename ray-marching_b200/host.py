"""ctypes view of librm_host.so — the C++ host that mirrors the reference's Rust API
(include/rm_host.h; ray-marching_b200/host/rm_host.hpp).  Used by tests and bench.py to build scenes,
cameras and uniforms through the product's own host code."""
from __future__ import annotations

import ctypes as C
from pathlib import Path

import numpy as np

HOST_LIB_PATH = Path(__file__).resolve().parent / "librm_host.so"
NONE, E_INVALID, E_PANIC, E_DEVICE = -1, -2, -3, -4
SPHERE, BOX, UNION, SUBTRACTION, INTERSECTION, SMOOTH_UNION, PLANE, TRANSLATION, ROTATION, SCALE = range(10)
VT_SCALAR, VT_VEC3, VT_CSGNODE = range(3)
EV_PAN, EV_ORBIT, EV_DOLLY = range(3)

HOST_SYMBOLS = [
    "rmh_reset", "rmh_last_error", "rmh_node_sphere", "rmh_node_box", "rmh_node_binary", "rmh_node_plane", "rmh_node_translation",
    "rmh_node_rotation", "rmh_node_rotation_matrix", "rmh_node_scale", "rmh_build_commands",
    "rmh_template_name", "rmh_template_param_count", "rmh_template_param", "rmh_template_evaluate", "rmh_orbit_new",
    "rmh_orbit_update", "rmh_orbit_state", "rmh_orbit_set", "rmh_orbit_camera", "rmh_camera_view", "rmh_uniforms",
    "rmh_limits_default", "rmh_graph_new", "rmh_graph_add_root", "rmh_graph_add_node", "rmh_graph_remove_node",
    "rmh_graph_set_scalar", "rmh_graph_set_vec3", "rmh_graph_connect", "rmh_graph_disconnect",
    "rmh_graph_evaluate_root", "rmh_graph_last_evaluations", "rmh_resources_new", "rmh_resources_ctx",
    "rmh_callback_render",
]

_lib = None


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        if not HOST_LIB_PATH.exists():
            raise RuntimeError(f"{HOST_LIB_PATH} is missing: run __graft_entry__.build() first")
        L = C.CDLL(str(HOST_LIB_PATH))
        L.rmh_last_error.restype = C.c_char_p
        L.rmh_template_name.restype = C.c_char_p
        L.rmh_template_param.restype = C.c_char_p
        L.rmh_resources_ctx.restype = C.c_void_p
        L.rmh_node_sphere.argtypes = [C.c_void_p, C.c_float]
        L.rmh_node_box.argtypes = [C.c_void_p, C.c_void_p]
        L.rmh_node_binary.argtypes = [C.c_int, C.c_int, C.c_int, C.c_float]
        L.rmh_node_plane.argtypes = [C.c_void_p, C.c_float]
        L.rmh_node_translation.argtypes = [C.c_int, C.c_void_p]
        L.rmh_node_rotation.argtypes = [C.c_int, C.c_void_p]
        L.rmh_node_rotation_matrix.argtypes = [C.c_int, C.c_void_p]
        L.rmh_node_scale.argtypes = [C.c_int, C.c_float]
        L.rmh_build_commands.argtypes = [C.c_int, C.c_void_p, C.c_uint32, C.c_void_p]
        L.rmh_template_param.argtypes = [C.c_int, C.c_int, C.c_void_p, C.c_void_p]
        L.rmh_template_evaluate.argtypes = [C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.rmh_orbit_new.argtypes = [C.c_void_p, C.c_float]
        L.rmh_orbit_update.argtypes = [C.c_int, C.c_int, C.c_float, C.c_float]
        L.rmh_orbit_state.argtypes = [C.c_int, C.c_void_p]
        L.rmh_orbit_set.argtypes = [C.c_int, C.c_float, C.c_float]
        L.rmh_orbit_camera.argtypes = [C.c_int, C.c_void_p, C.c_void_p]
        L.rmh_camera_view.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        L.rmh_uniforms.argtypes = [C.c_void_p, C.c_void_p, C.c_float, C.c_float, C.c_void_p]
        L.rmh_limits_default.argtypes = [C.c_void_p]
        L.rmh_graph_set_scalar.argtypes = [C.c_int, C.c_int, C.c_char_p, C.c_float]
        L.rmh_graph_set_vec3.argtypes = [C.c_int, C.c_int, C.c_char_p, C.c_void_p]
        L.rmh_graph_connect.argtypes = [C.c_int, C.c_int, C.c_int, C.c_char_p]
        L.rmh_graph_disconnect.argtypes = [C.c_int, C.c_int, C.c_char_p]
        L.rmh_callback_render.argtypes = [C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_float, C.c_float, C.c_void_p]
        _lib = L
    return _lib


def _f3(v):
    return np.ascontiguousarray(np.asarray(v, np.float32).reshape(3))


def sphere(center, radius) -> int:
    c = _f3(center)  # keep the arrays alive across the call (.ctypes.data is a bare address)
    return lib().rmh_node_sphere(c.ctypes.data, float(radius))


def box(center, radius) -> int:
    c, r = _f3(center), _f3(radius)
    return lib().rmh_node_box(c.ctypes.data, r.ctypes.data)


def binary(kind: int, a: int, b: int, k: float = 0.25) -> int:
    return lib().rmh_node_binary(kind, a, b, float(k))


def _checked(hnd: int, what: str) -> int:
    if hnd < 0:
        raise ValueError(f"{what}: {lib().rmh_last_error().decode() or hnd}")
    return hnd


# N4 node kinds (reserved in the reference at csg/mod.rs:34,41-44): one child, flattened as Push, child, Pop
def plane(normal, offset) -> int:
    n = _f3(normal)
    return _checked(lib().rmh_node_plane(n.ctypes.data, float(offset)), "plane")


def translation(child: int, offset) -> int:
    t = _f3(offset)
    return _checked(lib().rmh_node_translation(child, t.ctypes.data), "translation")


def rotation(child: int, rotation_wijk) -> int:
    q = np.ascontiguousarray(np.asarray(rotation_wijk, np.float32).reshape(4))
    return _checked(lib().rmh_node_rotation(child, q.ctypes.data), "rotation")


def rotation_matrix(child: int, m) -> int:
    a = np.ascontiguousarray(np.asarray(m, np.float32).reshape(9))
    return _checked(lib().rmh_node_rotation_matrix(child, a.ctypes.data), "rotation_matrix")


def scale(child: int, factor) -> int:
    return _checked(lib().rmh_node_scale(child, float(factor)), "scale")


def build_commands(node: int):
    """-> (cmd_count, words uint32[])  — BuildCommands::build_commands into a fresh builder."""
    cnt = C.c_uint32()
    n = lib().rmh_build_commands(node, None, 0, C.byref(cnt))
    if n < 0:
        raise ValueError("bad node handle")
    w = np.zeros(max(n, 1), np.uint32)
    lib().rmh_build_commands(node, w.ctypes.data, n, C.byref(cnt))
    return int(cnt.value), w[:n]


def scene_s1() -> int:
    return binary(SUBTRACTION, binary(UNION, box((0, 0, 0), (1, 1, 1)), sphere((0, 1, 0), 1.0)), sphere((0.6, 0.3, 1.0), 0.7))


class Orbit:
    """OrbitCameraController (camera.rs:21-85)."""

    def __init__(self, target=(0, 0, 0), radius=5.0):
        t = _f3(target)
        self.h = lib().rmh_orbit_new(t.ctypes.data, float(radius))

    def update(self, event: int, x: float, y: float = 0.0):
        lib().rmh_orbit_update(self.h, event, float(x), float(y))

    def set(self, yaw: float, pitch: float):
        lib().rmh_orbit_set(self.h, float(yaw), float(pitch))

    @property
    def state(self):
        o = np.zeros(7, np.float32)
        lib().rmh_orbit_state(self.h, o.ctypes.data)
        return {"target": o[:3].copy(), "pitch": float(o[3]), "yaw": float(o[4]), "radius": float(o[5])}

    def camera(self):
        p, q = np.zeros(3, np.float32), np.zeros(4, np.float32)
        lib().rmh_orbit_camera(self.h, p.ctypes.data, q.ctypes.data)
        return p, q

    def uniforms(self, W: float, H: float) -> bytes:
        p, q = self.camera()
        return uniforms(p, q, W, H)


def camera_view(position, rotation_wijk) -> np.ndarray:
    out = np.zeros(16, np.float32)
    p, q = _f3(position), np.ascontiguousarray(rotation_wijk, np.float32)
    lib().rmh_camera_view(p.ctypes.data, q.ctypes.data, out.ctypes.data)
    return out.reshape(4, 4).T  # column-major -> [row, col]


def uniforms(position, rotation_wijk, W: float, H: float) -> bytes:
    out = np.zeros(144, np.uint8)
    p, q = _f3(position), np.ascontiguousarray(rotation_wijk, np.float32)
    lib().rmh_uniforms(p.ctypes.data, q.ctypes.data, float(W), float(H), out.ctypes.data)
    return out.tobytes()


def limits_default() -> bytes:
    out = np.zeros(12, np.uint8)
    lib().rmh_limits_default(out.ctypes.data)
    return out.tobytes()


def template_params(kind: int):
    n = lib().rmh_template_param_count(kind)
    out = []
    for i in range(n):
        tag = C.c_int()
        v = np.zeros(3, np.float32)
        name = lib().rmh_template_param(kind, i, C.byref(tag), v.ctypes.data).decode()
        out.append((name, tag.value, v.copy()))
    return out


def template_evaluate(kind: int, params: dict) -> int:
    """params: name -> ("scalar", x) | ("vec3", xyz) | ("node", handle or NONE)"""
    names = (C.c_char_p * len(params))(*[k.encode() for k in params])
    tags = np.zeros(len(params), np.int32)
    vals = np.zeros((len(params), 3), np.float32)
    nodes = np.full(len(params), NONE, np.int32)
    for i, (t, v) in enumerate(params.values()):
        if t == "scalar":
            tags[i], vals[i, 0] = VT_SCALAR, v
        elif t == "vec3":
            tags[i], vals[i] = VT_VEC3, v
        else:
            tags[i], nodes[i] = VT_CSGNODE, v
    return lib().rmh_template_evaluate(kind, len(params), names, tags.ctypes.data, vals.ctypes.data, nodes.ctypes.data)


class Graph:
    """Headless CSGNodeGraph (csg_node_graph.rs:235-309)."""

    def __init__(self):
        self.h = lib().rmh_graph_new()

    def add_root(self):
        return lib().rmh_graph_add_root(self.h)

    def add_node(self, kind):
        return lib().rmh_graph_add_node(self.h, kind)

    def remove_node(self, n):
        return lib().rmh_graph_remove_node(self.h, n)

    def set_scalar(self, n, param, v):
        return lib().rmh_graph_set_scalar(self.h, n, param.encode(), float(v))

    def set_vec3(self, n, param, v):
        a = _f3(v)
        return lib().rmh_graph_set_vec3(self.h, n, param.encode(), a.ctypes.data)

    def connect(self, a, b, param):
        return lib().rmh_graph_connect(self.h, a, b, param.encode())

    def disconnect(self, b, param):
        return lib().rmh_graph_disconnect(self.h, b, param.encode())

    def evaluate_root(self):
        return lib().rmh_graph_evaluate_root(self.h)

    @property
    def evaluations(self):
        return lib().rmh_graph_last_evaluations(self.h)


def callback_render(res: int, node: int, position, rotation_wijk, W: int, H: int) -> np.ndarray:
    out = np.zeros((H, W, 4), np.uint8)
    p, q = _f3(position), np.ascontiguousarray(rotation_wijk, np.float32)
    rc = lib().rmh_callback_render(res, node, p.ctypes.data, q.ctypes.data, float(W), float(H), out.ctypes.data)
    if rc != 0:
        raise RuntimeError(f"rmh_callback_render rc={rc}: {lib().rmh_last_error().decode()}")
    return out
