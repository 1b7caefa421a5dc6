"""Known-answer tests K1–K6 pinning the CPU oracle to the shader text.

The reference has no tests or golden vectors of its own, so these analytically
derived answers are what pins oracle/rm_oracle.c to /root/reference/src/ray_marching/ray_marching.wgsl.
"""
import math

import numpy as np
import pytest

import helpers as h


def test_k1_centre_ray_hits_at_4():
    # wgsl:90-114,229-233 — step 0: d=4, step 1: d=0 < 0.01 -> hit at dist 4.0 exactly, 2 evals
    c, w = h.scene_s0()
    r = h.oracle_trace(c, w, [0, 0, 5], [0, 0, -1])
    assert r["cls"] == 2 and r["t"] == 4.0 and r["n_step"] == 2


def test_k2_centre_colour():
    # wgsl:100-105: l = normalize(pos-(2,-5,3)); n=(0,0,1); dot<0.02 -> 0.02 * albedo
    c, w = h.scene_s0()
    r = h.oracle_trace(c, w, [0, 0, 5], [0, 0, -1])
    np.testing.assert_allclose(r["rgb"], [0.008, 0.014, 0.002], rtol=1e-6)
    img = h.oracle_render(h.orbit_uniforms(640, 480), c, w, 640, 480)["rgba"]
    assert tuple(img[240, 320]) == (23, 30, 11, 255)


def test_k3_silhouette_radius():
    c, w = h.scene_s0()
    out = h.oracle_render(h.orbit_uniforms(640, 480), c, w, 640, 480, aov=True)
    hit = (out["cls"] == 2).any(-1)
    ys, xs = np.nonzero(hit)
    # effective radius r+min_dist = 1.01 -> 119.5 px (wgsl:97, renderer.rs:207)
    expect = 1.01 / math.sqrt(25 - 1.01 ** 2) / math.tan(math.pi / 8) * 240
    assert abs((xs.max() - xs.min() + 1) / 2 - expect) <= 1.0
    assert abs((ys.max() - ys.min() + 1) / 2 - expect) <= 1.0
    # centred
    assert xs.min() + xs.max() == 639 and ys.min() + ys.max() == 479


def test_k4_floor_and_sky_colours():
    c, w = h.scene_s0()
    out = h.oracle_render(h.orbit_uniforms(640, 480), c, w, 640, 480, aov=True)
    img, cls = out["rgba"], out["cls"]
    assert (img[..., 3] == 255).all()
    sky = (cls == 0).all(-1)
    assert sky.any() and (img[sky][:, :3] == 0).all()
    floor = (cls == 1).all(-1)
    cols = {tuple(p) for p in img[floor][:, :3]}
    assert (81, 81, 114) in cols and (140, 140, 161) in cols
    # top half is sky (horizon at image centre for pitch 0), bottom rows are floor
    assert sky[:100].all() and (cls[470:] == 1).all()


def test_k5_empty_scene_runs_all_iterations():
    # wgsl:188-191: cmd_count 0 -> map = max_dist; neither <0.01 nor >100 fires -> 100 steps
    r = h.oracle_trace(0, [], [0, 0, 5], [0, -0.6, -0.8])
    assert r["cls"] == 1 and r["n_step"] == 100
    out = h.oracle_render(h.orbit_uniforms(64, 48), 0, [], 64, 48, aov=True)
    assert not (out["cls"] == 2).any()
    assert out["counters"][0] == 64 * 48 * 16 * 100


def test_k6_inv_proj_w_is_one():
    # inv_proj * (x,y,-1,1) has w = 1/near = 1 (wgsl:57-59; renderer.rs:207)
    ip = np.frombuffer(h.orbit_uniforms(640, 480), np.float32)[4:20].reshape(4, 4).T
    for x, y in [(0.3, -0.7), (-1, 1), (0, 0)]:
        v = ip @ np.array([x, y, -1, 1], np.float32)
        assert abs(v[3] - 1.0) < 1e-6 and abs(v[2] + 1.0) < 1e-6


def test_uniform_layout_144_bytes():
    u = h.orbit_uniforms(640, 480)
    assert len(u) == 144
    f = np.frombuffer(u, np.float32)
    assert f[0] == 640 and f[1] == 480 and f[2] == 0 and f[3] == 0
    inv_view = f[20:36].reshape(4, 4).T
    np.testing.assert_allclose(inv_view[:, 3], [0, 0, 5, 1])
    np.testing.assert_allclose(inv_view[:3, :3], np.eye(3))


def test_q3_abort_compares_scene_distance():
    # wgsl:109: the abort test uses the scene distance, not travelled distance — a ray starting
    # > max_dist from everything aborts on step 0.
    c, w = h.scene_s0()
    r = h.oracle_trace(c, w, [0, 0, 200], [0, 0, -1])
    assert r["cls"] == 0 and r["n_step"] == 1


def test_q4_hit_at_distance_zero_inside_solid():
    c, w = h.scene_s0()
    r = h.oracle_trace(c, w, [0, 0, 0.5], [0, 0, -1])
    assert r["cls"] == 2 and r["t"] == 0.0 and r["n_step"] == 1


def test_q6_horizontal_ray_floor_is_defined():
    c, w = h.scene_s0()
    r = h.oracle_trace(c, w, [0, 3, 5], [1, 0, 0])  # direction.y == 0 -> -inf floor_dist
    assert r["cls"] == 0


def test_extension_ops_semantics():
    # X1 intersection = max(a,b); X2 smooth union polynomial
    a, b = h.sphere((0, 0, 0), 1.0), h.sphere((0.5, 0, 0), 1.0)
    p = np.array([[0.2, 0.1, 0.0], [2.0, 0, 0], [0.25, 0.9, 0.3]], np.float32)
    da = h.oracle_map(1, a, p)
    db = h.oracle_map(1, b, p)
    np.testing.assert_array_equal(h.oracle_map(3, a + b + [h.OP_INTERSECTION], p), np.maximum(da, db))
    np.testing.assert_array_equal(h.oracle_map(3, a + b + [h.OP_UNION], p), np.minimum(da, db))
    np.testing.assert_array_equal(h.oracle_map(3, a + b + [h.OP_SUBTRACTION], p), np.maximum(da, -db))
    k = np.float32(0.5)
    hh = np.maximum(k - np.abs(da - db), np.float32(0)) / k
    exp = np.minimum(da, db) - hh * hh * k * np.float32(0.25)
    np.testing.assert_array_equal(h.oracle_map(3, a + b + [h.OP_SMOOTH_UNION, h.fbits(k)], p), exp)


@pytest.mark.parametrize("words,count", [
    ([5], 1),                                  # unknown opcode (Q10)
    ([100], 1),                                # pops from empty stack
    ([0, 0, 0, 0], 1),                         # truncated sphere
    (h.sphere((0, 0, 0), 1) * 33, 33),         # depth 33 > 32 (wgsl:173)
])
def test_validation_rejects(words, count):
    assert h.oracle_validate(count, words) != 0


def test_k8_stack_depth_limits():
    s = h.sphere((0, 0, 0), 1)
    right_deep = s * 32 + [h.OP_UNION] * 31       # depth 32 = limit
    assert h.oracle_validate(63, right_deep) == 0
    left_deep = s + (s + [h.OP_UNION]) * 40       # depth 2, any length
    assert h.oracle_validate(81, left_deep) == 0
