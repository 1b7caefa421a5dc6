"""Host-side logic of the drop-in boundary, on CPU: the C++ mirror of the reference's Rust API
(CSG nodes + flattener, templates + ValueType, camera + orbit controller, uniform bytes, headless
graph evaluator) checked byte for byte against the independent numpy producers in helpers.py and
against the reference's documented behaviour (file:line in each test)."""
import math
import re
from pathlib import Path

import numpy as np
import pytest

import helpers as h
import rm_import

ROOT = Path(__file__).resolve().parent.parent


@pytest.fixture(scope="module")
def host():
    rm = rm_import.load()
    if not rm.HOST_LIB_PATH.exists():
        rm.build()
    import importlib
    return importlib.import_module("ray_marching_b200.host")


def test_host_library_exports_every_declared_symbol(host):
    header = (ROOT / "include" / "rm_host.h").read_text()
    declared = set(re.findall(r"RMH_API\s+[\w \*]+?\b(rmh_\w+)\s*\(", header))
    assert declared == set(host.HOST_SYMBOLS), declared ^ set(host.HOST_SYMBOLS)
    for s in declared:
        assert hasattr(host.lib(), s)


def test_k7_flatten_order_is_post_order(host):
    # operations/mod.rs:13-17: A, then B, then the operator => Subtraction(Union(A,B),C) = A B 100 C 101
    A = host.sphere((0, 0, 0), 1)
    B = host.box((1, 0, 0), (1, 2, 3))
    Cn = host.sphere((0, 2, 0), 0.5)
    cnt, w = host.build_commands(host.binary(host.SUBTRACTION, host.binary(host.UNION, A, B), Cn))
    assert cnt == 5
    expect = h.sphere((0, 0, 0), 1) + h.box((1, 0, 0), (1, 2, 3)) + [100] + h.sphere((0, 2, 0), 0.5) + [101]
    assert w.tolist() == expect


def test_wire_format_of_each_command(host):
    # builder.rs:39-61: opcode word first, then params as f32 bit patterns
    assert host.build_commands(host.sphere((1.5, -2, 3), 0.25))[1].tolist() == [0, h.fbits(1.5), h.fbits(-2), h.fbits(3), h.fbits(0.25)]
    assert host.build_commands(host.box((1, 2, 3), (4, 5, 6)))[1].tolist() == [1] + [h.fbits(x) for x in (1, 2, 3, 4, 5, 6)]
    a, b = host.sphere((0, 0, 0), 1), host.sphere((1, 0, 0), 1)
    assert host.build_commands(host.binary(host.INTERSECTION, a, b))[1][-1] == 102
    assert host.build_commands(host.binary(host.SMOOTH_UNION, a, b, 0.3))[1][-2:].tolist() == [103, h.fbits(0.3)]


def test_scenes_match_numpy_producers_byte_for_byte(host):
    cnt, w = host.build_commands(host.scene_s1())
    c1, w1 = h.scene_s1()
    assert cnt == c1 and np.array_equal(w, w1)
    # S2 through the C++ host, parameters drawn by the same SplitMix64 stream
    rng = h.SplitMix64(0x5EED0040)
    acc = None
    for i in range(32):
        cx, cy, cz = h._lerp(-2, 2, rng.u01()), h._lerp(-1, 1.5, rng.u01()), h._lerp(-2, 2, rng.u01())
        if i % 2 == 0:
            leaf = host.sphere((cx, cy, cz), h._lerp(0.25, 0.6, rng.u01()))
        else:
            leaf = host.box((cx, cy, cz), [h._lerp(0.2, 0.5, rng.u01()) for _ in range(3)])
        if acc is None:
            acc = leaf
        elif i % 8 == 0:
            acc = host.binary(host.SUBTRACTION, acc, leaf)
        else:
            acc = host.binary(host.SMOOTH_UNION, acc, leaf, 0.25)
    cnt, w = host.build_commands(acc)
    c2, w2 = h.scene_s2()
    assert cnt == c2 == 63 and np.array_equal(w, w2) and len(w) == 251


def test_template_names_params_and_defaults(host):
    # sphere.rs:26-43, box.rs:25-42, operations/mod.rs:25-42
    assert [host.lib().rmh_template_name(k).decode() for k in range(4)] == ["Sphere", "Box", "Union", "Subtraction"]
    p = host.template_params(host.SPHERE)
    assert [(n, t) for n, t, _ in p] == [("center", host.VT_VEC3), ("radius", host.VT_SCALAR)]
    assert p[0][2].tolist() == [0, 0, 0] and p[1][2][0] == 1.0
    p = host.template_params(host.BOX)
    assert [(n, t) for n, t, _ in p] == [("center", host.VT_VEC3), ("radius", host.VT_VEC3)] and p[1][2].tolist() == [1, 1, 1]
    for k in (host.UNION, host.SUBTRACTION):
        assert [(n, t) for n, t, _ in host.template_params(k)] == [("A", host.VT_CSGNODE), ("B", host.VT_CSGNODE)]


def test_template_evaluate(host):
    n = host.template_evaluate(host.SPHERE, {"center": ("vec3", (1, 2, 3)), "radius": ("scalar", 0.5)})
    assert host.build_commands(n)[1].tolist() == h.sphere((1, 2, 3), 0.5)
    a = host.sphere((0, 0, 0), 1)
    # operations/mod.rs:47-48: a missing connection (CSGNode(None)) yields None via `?`
    assert host.template_evaluate(host.UNION, {"A": ("node", a), "B": ("node", host.NONE)}) == host.NONE
    u = host.template_evaluate(host.UNION, {"A": ("node", a), "B": ("node", a)})
    assert host.build_commands(u)[0] == 3
    # sphere.rs:47-48: wrong value type / missing key unwrap()s -> panic
    assert host.template_evaluate(host.SPHERE, {"center": ("scalar", 1.0), "radius": ("scalar", 0.5)}) == host.E_PANIC
    assert host.template_evaluate(host.SPHERE, {"center": ("vec3", (0, 0, 0))}) == host.E_PANIC


def test_orbit_camera_matches_closed_form(host):
    # camera.rs:52-60 with from_euler_angles(-pitch, -yaw, 0): R = Ry(-yaw) Rx(-pitch), eye = target + R z r
    for yaw, pitch in [(0, 0), (0.6, 0.4), (-2.0, 1.2), (3.0, -0.7)]:
        o = host.Orbit((0.5, -0.25, 1.0), 5.0)
        o.set(yaw, pitch)
        pos, q = o.camera()
        eye, R = h.orbit_camera((0.5, -0.25, 1.0), 5.0, yaw, pitch)
        np.testing.assert_allclose(pos, eye, atol=2e-6)
        view = host.camera_view(pos, q)  # camera.rs:8-13: R^-1 * T(-pos)
        expect = np.eye(4)
        expect[:3, :3] = R.T
        expect[:3, 3] = -R.T @ eye
        np.testing.assert_allclose(view, expect, atol=3e-6)
    o = host.Orbit((0, 0, 0), 5.0)  # main.rs:38
    pos, q = o.camera()
    assert pos.tolist() == [0, 0, 5] and q.tolist() == [1, 0, 0, 0]


def test_uniform_bytes_layout_and_values(host):
    # renderer.rs:205-222 + WGSL uniform layout (extent @0, inv_proj @16, inv_view @80)
    for (W, H, yaw, pitch) in [(640, 480, 0, 0), (1920, 1080, 0.6, 0.4), (3840, 2160, 2.5, -1.0)]:
        o = host.Orbit((0, 0, 0), 5.0)
        o.set(yaw, pitch)
        got = np.frombuffer(o.uniforms(W, H), np.float32)
        ref = np.frombuffer(h.orbit_uniforms(W, H, yaw=yaw, pitch=pitch), np.float32)
        assert got[0] == W and got[1] == H and got[2] == 0 and got[3] == 0
        np.testing.assert_allclose(got[4:20], ref[4:20], rtol=1e-6, atol=1e-9)
        np.testing.assert_allclose(got[20:36], ref[20:36], atol=3e-6)
        # K6: w of inv_proj * (x, y, -1, 1) is 1/near = 1
        ip = got[4:20].reshape(4, 4).T
        assert abs((ip @ np.array([0.3, -0.2, -1, 1], np.float32))[3] - 1.0) < 1e-6
    assert host.limits_default() == h.limits_bytes(0.01, 100.0, 100)  # renderer.rs:133-137


def test_orbit_controller_update_events(host):
    o = host.Orbit((0, 0, 0), 5.0)
    o.update(host.EV_ORBIT, 30.0, 20.0)  # camera.rs:71-73: += d * 0.01
    assert o.state["yaw"] == pytest.approx(0.3) and o.state["pitch"] == pytest.approx(0.2)
    o.update(host.EV_ORBIT, 0.0, 1000.0)  # camera.rs:77: pitch clamped to +-1.5
    assert o.state["pitch"] == pytest.approx(1.5)
    o.update(host.EV_ORBIT, 0.0, -5000.0)
    assert o.state["pitch"] == pytest.approx(-1.5)
    o = host.Orbit((0, 0, 0), 5.0)
    o.update(host.EV_DOLLY, 10.0)  # camera.rs:80: radius += delta * 0.01 * radius
    assert o.state["radius"] == pytest.approx(5.5)
    o.update(host.EV_DOLLY, -1000.0)  # camera.rs:81: radius >= 0.1
    assert o.state["radius"] == pytest.approx(0.1)
    o = host.Orbit((0, 0, 0), 5.0)
    o.update(host.EV_PAN, 100.0, 50.0)  # camera.rs:64-69 at identity rotation: target += (-dx, +dy, 0) * 0.01
    np.testing.assert_allclose(o.state["target"], [-1.0, 0.5, 0.0], atol=1e-6)
    o = host.Orbit((0, 0, 0), 5.0)
    o.set(math.pi / 2, 0.0)  # yaw 90 deg: rotation = Ry(-yaw); right = R x = (0, 0, 1)
    o.update(host.EV_PAN, 100.0, 0.0)
    np.testing.assert_allclose(o.state["target"], [0.0, 0.0, -1.0], atol=1e-6)


def test_graph_evaluate_root(host):
    g = host.Graph()
    assert g.evaluate_root() == host.NONE  # no Root node (csg_node_graph.rs:252-257)
    root = g.add_root()
    assert g.evaluate_root() == host.NONE  # unconnected "SDF" input -> CSGNode(None)
    bx = g.add_node(host.BOX)
    sp = g.add_node(host.SPHERE)
    g.set_vec3(sp, "center", (0, 1, 0))
    un = g.add_node(host.UNION)
    g.connect(bx, un, "A")
    g.connect(un, root, "SDF")
    assert g.evaluate_root() == host.NONE  # Union input B missing -> None propagates (operations/mod.rs:47-48)
    g.connect(sp, un, "B")
    s2 = g.add_node(host.SPHERE)
    g.set_vec3(s2, "center", (0.6, 0.3, 1.0))
    g.set_scalar(s2, "radius", 0.7)
    sub = g.add_node(host.SUBTRACTION)
    g.connect(un, sub, "A")
    g.connect(s2, sub, "B")
    g.connect(sub, root, "SDF")
    cnt, w = host.build_commands(g.evaluate_root())
    c1, w1 = h.scene_s1()
    assert cnt == c1 and np.array_equal(w, w1)  # template defaults: box c=0 r=1, sphere r=1
    g.remove_node(s2)
    assert g.evaluate_root() == host.NONE


def test_graph_connect_refuses_a_data_type_mismatch(host):
    # csg_node_graph.rs:17-22,84-101: the editor only joins ports of equal DataType; every node output is an SDF
    g = host.Graph()
    root = g.add_root()
    sp, bx, un = g.add_node(host.SPHERE), g.add_node(host.BOX), g.add_node(host.UNION)
    assert g.connect(bx, sp, "center") == host.E_INVALID    # SDF -> Vec3
    assert g.connect(bx, sp, "radius") == host.E_INVALID    # SDF -> Scalar
    assert g.connect(sp, bx, "radius") == host.E_INVALID    # SDF -> Vec3 (box radius)
    sm = g.add_node(host.SMOOTH_UNION)
    assert g.connect(sp, sm, "k") == host.E_INVALID
    assert g.connect(sp, sm, "A") == 0 and g.connect(bx, sm, "B") == 0 and g.connect(sm, root, "SDF") == 0
    # the refused connections left the inputs as they were: the tree is SmoothUnion(Sphere default, Box default)
    cnt, w = host.build_commands(g.evaluate_root())
    assert cnt == 3 and w.tolist() == h.sphere((0, 0, 0), 1.0) + h.box((0, 0, 0), (1, 1, 1)) + [h.OP_SMOOTH_UNION, w.tolist()[-1]]
    assert g.connect(un, un, "nope") == host.E_PANIC         # unknown port: still where the reference would panic


def test_graph_evaluator_memoises_shared_nodes(host):
    # csg_node_graph.rs:284-289: output_cache — a node feeding two inputs is evaluated once
    g = host.Graph()
    root = g.add_root()
    sp = g.add_node(host.SPHERE)
    u1 = g.add_node(host.UNION)
    g.connect(sp, u1, "A")
    g.connect(sp, u1, "B")
    u2 = g.add_node(host.UNION)
    g.connect(u1, u2, "A")
    g.connect(u1, u2, "B")
    g.connect(u2, root, "SDF")
    n = g.evaluate_root()
    assert host.build_commands(n)[0] == 7  # 4 sphere leaves + 3 unions after tree expansion
    assert g.evaluations == 3              # but only sp, u1, u2 were evaluated
    g.connect(u2, u1, "B")                 # a cycle: the reference would recurse forever; reported as a panic
    assert g.evaluate_root() == host.E_PANIC


# ---- N4: the node kinds the reference reserves in comments (csg/mod.rs:34,41-44; builder.rs:8,16-23) ----------------
def test_n4_nodes_flatten_to_push_child_pop(host):
    s = host.sphere((0, 0, 0), 1.0)
    cnt, w = host.build_commands(host.translation(s, (1, 2, 3)))
    assert cnt == 3 and w.tolist() == h.translated((1, 2, 3), (h.sphere((0, 0, 0), 1.0), 1))[0]
    cnt, w = host.build_commands(host.scale(s, 2.5))
    assert cnt == 3 and w.tolist() == h.scaled(2.5, (h.sphere((0, 0, 0), 1.0), 1))[0]
    R = [[0, -1, 0], [1, 0, 0], [0, 0, 1]]
    cnt, w = host.build_commands(host.rotation_matrix(s, R))
    assert cnt == 3 and w.tolist() == h.rotated(R, (h.sphere((0, 0, 0), 1.0), 1))[0]
    # the plane's normal is normalised by the constructor; opcode first, then normal, then offset
    cnt, w = host.build_commands(host.plane((0, 2, 0), 1.5))
    assert cnt == 1 and w.tolist() == h.plane((0, 1, 0), 1.5)
    n = np.array([1, 2, 2], np.float32)
    w = host.build_commands(host.plane(n, -0.5))[1]
    np.testing.assert_allclose(w[1:4].view(np.float32), n / 3, rtol=1e-6)


def test_n4_rotation_from_quaternion_is_the_same_rotation_the_camera_uses(host):
    # a quarter turn about z as a unit quaternion (w, i, j, k) -> row-major matrix, object -> world
    q = (math.cos(math.pi / 4), 0, 0, math.sin(math.pi / 4))
    w = host.build_commands(host.rotation(host.sphere((0, 0, 0), 1.0), q))[1]
    m = w[1:10].view(np.float32).reshape(3, 3)
    np.testing.assert_allclose(m, [[0, -1, 0], [1, 0, 0], [0, 0, 1]], atol=1e-6)
    # and for a general axis it agrees with Rodrigues' formula (helpers.rotation_matrix)
    axis, ang = np.array([1.0, 2.0, 3.0]), 0.7
    a = axis / np.linalg.norm(axis)
    q = (math.cos(ang / 2), *(a * math.sin(ang / 2)))
    w = host.build_commands(host.rotation(host.sphere((0, 0, 0), 1.0), q))[1]
    np.testing.assert_allclose(w[1:10].view(np.float32).reshape(3, 3), h.rotation_matrix(axis, ang), atol=1e-6)


def test_n4_nested_tree_is_valid_and_means_what_it_says(host):
    # Scale(2, Rotation(Translation(Union(sphere, box)))) smooth-united with a ground plane
    inner = host.binary(host.UNION, host.sphere((0, 0, 0), 0.5), host.box((0.6, 0, 0), (0.3, 0.3, 0.3)))
    R = h.rotation_matrix((0, 0, 1), 0.5)
    node = host.scale(host.rotation_matrix(host.translation(inner, (1, 0, 0)), R), 2.0)
    scene = host.binary(host.SMOOTH_UNION, node, host.plane((0, 1, 0), 1.5), 0.25)
    cnt, w = host.build_commands(scene)
    assert cnt == 11 and h.oracle_validate(cnt, w) == 0
    exp = h.scaled(2.0, h.rotated(R, h.translated((1, 0, 0), (h.sphere((0, 0, 0), 0.5) + h.box((0.6, 0, 0), (0.3, 0.3, 0.3)) + [100], 3))))
    assert w.tolist() == exp[0] + h.plane((0, 1, 0), 1.5) + [103, h.fbits(0.25)]
    # the centre of the inner sphere ends up at 2 * R (1,0,0) in the world
    c = 2.0 * (R @ np.array([1.0, 0, 0]))
    d = h.oracle_map(cnt, w, np.array([c], np.float32))
    assert abs(d[0] - (-1.0)) < 1e-5  # radius 0.5 scaled by 2


def test_n4_constructors_reject_what_the_validator_would(host):
    s = host.sphere((0, 0, 0), 1.0)
    for bad in (0.0, -1.0, float("inf"), float("nan")):
        with pytest.raises(ValueError):
            host.scale(s, bad)
    with pytest.raises(ValueError):
        host.plane((0, 0, 0), 1.0)
    with pytest.raises(ValueError):
        host.translation(12345, (0, 0, 0))  # not a node handle


def test_n4_templates_and_graph(host):
    names = [host.lib().rmh_template_name(k).decode() for k in range(10)]
    assert names[6:] == ["Plane", "Translation", "Rotation", "Scale"]
    assert [(n, t) for n, t, _ in host.template_params(host.PLANE)] == [("normal", host.VT_VEC3), ("offset", host.VT_SCALAR)]
    assert [(n, t) for n, t, _ in host.template_params(host.TRANSLATION)] == [("A", host.VT_CSGNODE), ("offset", host.VT_VEC3)]
    assert [(n, t) for n, t, _ in host.template_params(host.ROTATION)] == [("A", host.VT_CSGNODE), ("axis", host.VT_VEC3), ("angle", host.VT_SCALAR)]
    p = host.template_params(host.SCALE)
    assert [(n, t) for n, t, _ in p] == [("A", host.VT_CSGNODE), ("factor", host.VT_SCALAR)] and p[1][2][0] == 1.0
    # a missing child propagates None like the binary operators' `?` (operations/mod.rs:47-48)
    assert host.template_evaluate(host.SCALE, {"A": ("node", host.NONE), "factor": ("scalar", 2.0)}) == host.NONE
    assert host.template_evaluate(host.SCALE, {"A": ("node", host.sphere((0, 0, 0), 1)), "factor": ("scalar", -2.0)}) == host.E_PANIC
    # the editor graph: Sphere -> Translation -> Rotation(defaults: angle 0 = identity) -> Root
    g = host.Graph()
    root = g.add_root()
    sp = g.add_node(host.SPHERE)
    tr = g.add_node(host.TRANSLATION)
    g.set_vec3(tr, "offset", (0, 1, 0))
    ro = g.add_node(host.ROTATION)
    g.connect(sp, tr, "A")
    g.connect(tr, ro, "A")
    g.connect(ro, root, "SDF")
    cnt, w = host.build_commands(g.evaluate_root())
    assert cnt == 5 and w[0] == 202 and w[10] == 200 and w[-2:].tolist() == [201, 203]
    np.testing.assert_array_equal(w[1:10].view(np.float32), np.eye(3, dtype=np.float32).reshape(9))
    pts = np.random.default_rng(0).uniform(-3, 3, (100, 3)).astype(np.float32)
    np.testing.assert_allclose(h.oracle_map(cnt, w, pts), h.oracle_map(1, h.sphere((0, 1, 0), 1.0), pts), atol=1e-6)
    g.disconnect(tr, "A")
    assert g.evaluate_root() == host.NONE
