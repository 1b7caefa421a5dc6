"""Known-answer tests of the N4 extension on the CPU oracle: Plane (opcode 2) and the
space-transform groups TranslationPush/Pop (200/201), RotationPush/Pop (202/203), ScalePush/Pop (204/205) —
the opcode values /root/reference/src/ray_marching/csg/builder.rs:8,16-23 reserves in comments; the reference
itself implements none of them, so the semantics are build-defined (oracle/rm_oracle.c map_scene) and pinned
here by cases whose answers follow from the commands the reference does implement.
"""
import numpy as np
import pytest

import helpers as h
from oracle import np_oracle

PTS = np.random.default_rng(5).uniform(-4, 4, (2000, 3)).astype(np.float32)


def test_translation_moves_the_child():
    # (p - t) - 0 == p - c for c = t: the translated unit sphere IS the sphere centred at t, bit for bit
    t = (1.0, -0.5, 0.25)
    w, c = h.translated(t, (h.sphere((0, 0, 0), 1.0), 1))
    assert c == 3
    np.testing.assert_array_equal(h.oracle_map(c, w, PTS), h.oracle_map(1, h.sphere(t, 1.0), PTS))


def test_rotation_by_a_quarter_turn_permutes_the_box_radii():
    # R = Rz(90 deg) maps object x to world y; R^T p = (py, -px, pz), and |.| drops the sign
    R = [[0, -1, 0], [1, 0, 0], [0, 0, 1]]
    w, c = h.rotated(R, (h.box((0, 0, 0), (1.0, 2.0, 0.5)), 1))
    np.testing.assert_array_equal(h.oracle_map(c, w, PTS), h.oracle_map(1, h.box((0, 0, 0), (2.0, 1.0, 0.5)), PTS))


def test_rotation_uses_the_transpose():
    # a box centred at (1,0,0) in object space, turned +90 deg about z, sits at (0,1,0) in the world
    R = [[0, -1, 0], [1, 0, 0], [0, 0, 1]]
    w, c = h.rotated(R, (h.box((1, 0, 0), (0.2, 0.2, 0.2)), 1))
    d = h.oracle_map(c, w, np.array([[0, 1, 0], [1, 0, 0], [0, -1, 0]], np.float32))
    assert d[0] == np.float32(-0.2) and d[1] > 0.9 and d[2] > 1.7


def test_scale_by_a_power_of_two_is_exact():
    w, c = h.scaled(2.0, (h.sphere((0, 0, 0), 1.0), 1))
    np.testing.assert_array_equal(h.oracle_map(c, w, PTS), h.oracle_map(1, h.sphere((0, 0, 0), 2.0), PTS))
    w, c = h.scaled(0.5, (h.box((0, 0, 0), (1.0, 2.0, 4.0)), 1))
    np.testing.assert_array_equal(h.oracle_map(c, w, PTS), h.oracle_map(1, h.box((0, 0, 0), (0.5, 1.0, 2.0)), PTS))


def test_plane_is_dot_plus_offset():
    w = h.plane((0, 1, 0), 1.5)
    np.testing.assert_array_equal(h.oracle_map(1, w, PTS), PTS[:, 1] + np.float32(1.5))
    n = np.array([0.6, 0.0, 0.8], np.float32)
    exp = ((PTS[:, 0] * n[0] + PTS[:, 1] * n[1]) + PTS[:, 2] * n[2]) + np.float32(-0.25)
    np.testing.assert_array_equal(h.oracle_map(1, h.plane(n, -0.25), PTS), exp)


def test_pop_restores_the_evaluation_point():
    a = h.translated((3, 0, 0), (h.sphere((0, 0, 0), 1.0), 1))
    b = (h.box((0, 0, 0), (1, 1, 1)), 1)
    w = a[0] + b[0] + [h.OP_UNION]
    exp = np.minimum(h.oracle_map(1, h.sphere((3, 0, 0), 1.0), PTS), h.oracle_map(1, b[0], PTS))
    np.testing.assert_array_equal(h.oracle_map(a[1] + 2, w, PTS), exp)


def test_nested_groups_compose_outermost_first():
    # Translate(t, Scale(2, sphere r=1 at c)) == sphere r=2 at t + 2c, up to the two extra roundings of the chain
    inner = h.scaled(2.0, (h.sphere((0.5, 0, -1), 1.0), 1))
    w, c = h.translated((1, 2, 0), inner)
    np.testing.assert_allclose(h.oracle_map(c, w, PTS), h.oracle_map(1, h.sphere((2, 2, -2), 2.0), PTS), rtol=0, atol=2e-6)
    # the other nesting order scales the translation too: Scale(2, Translate(t, .)) moves by 2t
    w, c = h.scaled(2.0, h.translated((1, 2, 0), (h.sphere((0, 0, 0), 1.0), 1)))
    np.testing.assert_allclose(h.oracle_map(c, w, PTS), h.oracle_map(1, h.sphere((2, 4, 0), 2.0), PTS), rtol=0, atol=2e-6)


def test_scale_pop_rescales_the_whole_child_not_each_leaf():
    # s * smooth(a, b) with a, b evaluated at p / s
    k = np.float32(0.25)
    child = (h.sphere((0, 0, 0), 1.0) + h.box((0.75, 0, 0), (0.5, 0.5, 0.5)) + [h.OP_SMOOTH_UNION, h.fbits(k)], 3)
    w, c = h.scaled(3.0, child)
    q = (PTS / np.float32(3.0)).astype(np.float32)
    exp = (h.oracle_map(3, child[0], q) * np.float32(3.0)).astype(np.float32)
    np.testing.assert_array_equal(h.oracle_map(c, w, PTS), exp)


S = h.sphere((0, 0, 0), 1)


@pytest.mark.parametrize("words,count", [
    ([h.OP_TRANSLATION_POP], 1),                                                    # pop without push
    ([h.OP_TRANSLATION_PUSH, 0, 0, 0] + S, 2),                                      # never closed
    ([h.OP_TRANSLATION_PUSH, 0, 0, 0] + S + [h.OP_ROTATION_POP], 3),                # closed by the wrong kind
    ([h.OP_TRANSLATION_PUSH, 0, 0, 0] + [h.OP_TRANSLATION_POP], 2),                 # group without a child
    ([h.OP_TRANSLATION_PUSH, 0, 0, 0] + S + S + [h.OP_TRANSLATION_POP, h.OP_UNION], 5),  # two children
    (S + [h.OP_TRANSLATION_PUSH, 0, 0, 0] + S + [h.OP_UNION, h.OP_TRANSLATION_POP], 5),  # operator reaches outside
    ([h.OP_SCALE_PUSH, h.fbits(2.0)] + S + [h.OP_SCALE_POP, h.fbits(3.0)], 3),      # pop with another factor
    ([h.OP_SCALE_PUSH, h.fbits(0.0)] + S + [h.OP_SCALE_POP, h.fbits(0.0)], 3),      # s must be > 0
    ([h.OP_SCALE_PUSH, h.fbits(-1.0)] + S + [h.OP_SCALE_POP, h.fbits(-1.0)], 3),
    ([h.OP_SCALE_PUSH, h.fbits(float("inf"))] + S + [h.OP_SCALE_POP, h.fbits(float("inf"))], 3),
    ([h.OP_ROTATION_PUSH] + [0] * 8, 1),                                            # truncated matrix
    ([h.OP_TRANSLATION_PUSH, 0, 0, 0] * 9 + S + [h.OP_TRANSLATION_POP] * 9, 19),    # nesting deeper than 8
])
def test_validation_rejects_malformed_groups(words, count):
    assert h.oracle_validate(count, words) != 0


def test_validation_accepts_nesting_up_to_8():
    w = [h.OP_TRANSLATION_PUSH, 0, 0, 0] * 8 + S + [h.OP_TRANSLATION_POP] * 8
    assert h.oracle_validate(17, w) == 0


@pytest.mark.parametrize("seed", range(8))
def test_numpy_restatement_agrees_bit_for_bit_on_random_extended_programs(seed):
    rng = np.random.default_rng(100 + seed)
    n_ext = 0
    for n_leaves in (1, 2, 3, 5, 9, 14):
        words, cmd = h.random_program(rng, n_leaves, ext=True)
        assert h.oracle_validate(cmd, words) == 0
        n_ext += any(op in words for op in (200, 202, 204))
        ref = h.oracle_map(cmd, words, PTS)
        got = np_oracle.map_scene(cmd, np.array(words, np.uint32), PTS[:, 0].copy(), PTS[:, 1].copy(), PTS[:, 2].copy())
        assert np.array_equal(got.view(np.uint32), ref.view(np.uint32))
    assert n_ext >= 2


def test_rendered_frame_agrees_between_the_two_restatements():
    rng = np.random.default_rng(3)
    words, cmd = h.random_program(rng, 6, ext=True)
    ground = h.plane((0, 1, 0), 1.0)
    words, cmd = words + ground + [h.OP_SMOOTH_UNION, h.fbits(0.5)], cmd + 2
    u = h.orbit_uniforms(40, 24, yaw=0.6, pitch=0.4)
    a = h.oracle_render(u, cmd, words, 40, 24, aov=True)
    b = np_oracle.render(u, cmd, np.array(words, np.uint32), 40, 24)
    assert np.array_equal(a["rgba"], b["rgba"])
    assert np.array_equal(a["cls"], b["cls"])
    assert np.array_equal(a["t"].view(np.uint32), b["t"].view(np.uint32))
    assert np.array_equal(a["counters"], b["counters"])
    assert (a["cls"] == 2).mean() > 0.2


# ---- golden frames of the N4 scene S3 (tests/golden/n4, made by oracle/make_golden.py) -------------------------------
from pathlib import Path  # noqa: E402

GOLDEN_N4 = sorted((Path(__file__).resolve().parent / "golden" / "n4").glob("*.npz"))


@pytest.mark.parametrize("path", GOLDEN_N4, ids=lambda p: p.stem)
def test_n4_golden_frames_are_reproduced_by_both_restatements(path):
    z = np.load(path)
    g = {k: z[k] for k in z.files}
    lim = g["limits"].tobytes()
    W, H, cmd = int(g["W"]), int(g["H"]), int(g["cmd_count"])
    r = h.oracle_render(g["uniforms"].tobytes(), cmd, g["words"], W, H, limits=lim, aov=True)
    assert np.array_equal(r["rgba"], g["rgba"]) and np.array_equal(r["cls"], g["cls"])
    assert np.array_equal(r["t"].view(np.uint32), g["t"].view(np.uint32))
    assert np.array_equal(r["counters"], g["counters"])
    limits = (*np.frombuffer(lim[:8], np.float32), int(np.frombuffer(lim[8:], np.uint32)[0]))
    n = np_oracle.render(g["uniforms"].tobytes(), cmd, g["words"], W, H, limits)
    assert np.array_equal(n["rgba"], g["rgba"]) and np.array_equal(n["cls"], g["cls"])
    hit = g["cls"] == 2
    assert np.array_equal(n["t"][hit], g["t"][hit])
    assert np.array_equal(n["counters"], g["counters"])


def test_s3_definition():
    cmd, words = h.scene_s3()
    assert cmd == 65 and len(words) == 244 and h.oracle_validate(cmd, words) == 0
    # the first instance (i = 0: identity rotation, translation (1.6, -0.3, 0)) is the box centred there
    d = h.oracle_map(cmd, words, np.array([[1.6, -0.3, 0.0]], np.float32))
    assert abs(d[0] - (-0.2)) < 1e-5
