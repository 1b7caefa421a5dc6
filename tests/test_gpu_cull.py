"""Opt-in per-tile culling of the program (rm_set_cull, csrc/rm_cull.cu): every block interprets only the instructions that
can change the accumulator somewhere in its tile's cone.  The claim is EXACTNESS: the frame and the work counters are
identical to the unculled paint — and, in exact mode, to the oracle — while fewer instructions run per evaluation."""
import numpy as np
import pytest

import helpers as h
import rm_import
from test_gpu_parity import random_program

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def rm():
    return rm_import.load()


@pytest.fixture()
def ctx(rm):
    c = rm.Context(0)
    yield c
    c.close()


def on_off(ctx, u, cmd, words, W, H, **kw):
    out = []
    for on in (False, True):
        ctx.set_cull(on)
        ctx.set_counters(True)
        img = ctx.render(u, cmd, words, W, H, **kw)
        out.append((img, ctx.get_counters()))
        ctx.set_counters(False)
    ctx.set_cull(False)
    return out


@pytest.mark.parametrize("name,W,H,yaw,pitch,radius", [("S2", 384, 216, 0.6, 0.4, 5.0), ("S2", 300, 170, 2.4, 0.1, 3.0),
                                                       ("S2", 256, 144, 1.1, 0.7, 1.2), ("S2", 131, 77, 4.0, -0.3, 0.4)])
@pytest.mark.parametrize("mode", ["exact", "fast"])
def test_culled_equals_unculled_bit_for_bit(rm, ctx, name, W, H, yaw, pitch, radius, mode):
    cmd, words = h.SCENES[name]()
    u = h.orbit_uniforms(W, H, yaw=yaw, pitch=pitch, radius=radius)
    ctx.set_mode(rm.MODE_EXACT if mode == "exact" else rm.MODE_FAST)
    ctx.cull_stats()
    (a, ca), (b, cb) = on_off(ctx, u, cmd, words, W, H)
    st = ctx.cull_stats()
    assert st["culled_paints"] == 1
    assert np.array_equal(a, b), int((a != b).any(-1).sum())
    assert np.array_equal(ca, cb), (ca, cb)
    if mode == "exact":
        ref = h.oracle_render(u, cmd, words, W, H)
        assert np.array_equal(b, ref["rgba"]) and np.array_equal(cb, ref["counters"])
    # never more work than the full program (how much less depends on how narrow the tiles' cones are: at these sizes a
    # pixel is ten times a 4K pixel and little can be dropped; the 3840x2160 test below asks for a real reduction)
    if name == "S2":
        evals = int(cb[0]) + 4 * int(cb[1])
        assert 0 < st["executed_map_flops"] <= 798 * evals, (st, evals)


@pytest.mark.parametrize("seed", range(6))
def test_random_trees_culled_equals_oracle(rm, ctx, seed):
    """mixed tree shapes (push / pop instructions are always kept), all four operators, several k"""
    rng = np.random.default_rng(4242 + seed)
    W, H = 88, 64
    u = h.orbit_uniforms(W, H, yaw=0.7 * seed, pitch=0.25, radius=4.0)
    ctx.set_mode(rm.MODE_EXACT)
    for n_leaves in (1, 2, 5, 11, 24):
        words, cmd = random_program(rng, n_leaves)
        if len(words) > 255:
            continue
        (a, ca), (b, cb) = on_off(ctx, u, cmd, words, W, H)
        ref = h.oracle_render(u, cmd, words, W, H)
        assert np.array_equal(b, ref["rgba"]) and np.array_equal(cb, ref["counters"]), n_leaves
        assert np.array_equal(a, b)


@pytest.mark.parametrize("k", [0.05, 0.25, 1.0, 3.0])
def test_wide_and_narrow_smooth_unions(rm, ctx, k):
    """the blend reaches k away from a leaf: a wide k must keep far leaves alive"""
    W, H = 160, 90
    u = h.orbit_uniforms(W, H, yaw=0.3, pitch=0.3, radius=6.0)
    words = h.sphere((-1.5, 0, 0), 0.6)
    for i in range(1, 7):
        leaf = h.box((-1.5 + 0.6 * i, 0.2 * (i % 3), 0.1 * i), (0.3, 0.25, 0.2)) if i % 2 else h.sphere((-1.5 + 0.6 * i, 0, 0), 0.4)
        words = words + leaf + [h.OP_SMOOTH_UNION, h.fbits(k)]
    words = words + h.sphere((0, 0.5, 0.6), 0.5) + [h.OP_SUBTRACTION] + h.box((0, 0, 0), (3, 0.8, 0.8)) + [h.OP_INTERSECTION]
    cmd = 1 + 2 * 6 + 4
    for mode in (rm.MODE_EXACT, rm.MODE_FAST):
        ctx.set_mode(mode)
        (a, ca), (b, cb) = on_off(ctx, u, cmd, words, W, H)
        assert np.array_equal(a, b) and np.array_equal(ca, cb), (k, mode)
    ctx.set_mode(rm.MODE_EXACT)
    ctx.set_cull(True)
    assert np.array_equal(ctx.render(u, cmd, words, W, H), h.oracle_render(u, cmd, words, W, H)["rgba"])


def test_bands_strips_ragged_sizes_limits_and_extent(rm, ctx):
    import importlib
    sh = importlib.import_module("ray_marching_b200.sharding")
    cmd, words = h.scene_s2()
    ctx.set_mode(rm.MODE_EXACT)
    for W, H in ((37, 23), (9, 5), (130, 3)):
        u = h.orbit_uniforms(W, H, yaw=0.6, pitch=0.4)
        (a, ca), (b, cb) = on_off(ctx, u, cmd, words, W, H)
        assert np.array_equal(a, b) and np.array_equal(ca, cb), (W, H)
    W, H, strip, world = 160, 50, 16, 3
    u = h.orbit_uniforms(W, H, yaw=0.6, pitch=0.4)
    ctx.set_cull(False)
    full = ctx.render(u, cmd, words, W, H)
    ctx.set_cull(True)
    assert np.array_equal(ctx.render(u, cmd, words, W, H, row0=17, nrows=20), full[17:37])
    rows_max = sh.max_local_rows(H, strip, world)
    gathered = np.zeros((world, rows_max, W, 4), np.uint8)
    ctx.prepare(u, cmd, words)
    for r in range(world):
        n = ctx.paint_strips(W, H, strip, r, world)
        if n:
            ctx.download(gathered[r, :n])
    assert np.array_equal(sh.reassemble(gathered, H, strip, world), full)
    # run-time limits: a small max_dist and a tiny iteration budget change where rays stop, not the identity
    for lim in ((0.01, 3.0, 100), (0.05, 100.0, 7), (0.001, 40.0, 300)):
        ctx.set_limits(*lim)
        (a, ca), (b, cb) = on_off(ctx, u, cmd, words, W, H)
        assert np.array_equal(a, b) and np.array_equal(ca, cb), lim
        assert np.array_equal(b, h.oracle_render(u, cmd, words, W, H, limits=h.limits_bytes(*lim))["rgba"]), lim
    ctx.set_limits(0.01, 100.0, 100)
    # a viewport_extent that is not the frame size spreads the samples differently (wgsl:53)
    uu = bytearray(u)
    uu[0:8] = np.array([W / 3.0, H * 1.7], np.float32).tobytes()
    (a, ca), (b, cb) = on_off(ctx, bytes(uu), cmd, words, W, H)
    assert np.array_equal(a, b) and np.array_equal(ca, cb)
    assert np.array_equal(b, h.oracle_render(bytes(uu), cmd, words, W, H)["rgba"])


def test_programs_it_does_not_apply_to_paint_as_usual(rm, ctx):
    W, H = 64, 36
    u = h.orbit_uniforms(W, H, yaw=0.6, pitch=0.4)
    ctx.set_mode(rm.MODE_EXACT)
    ctx.set_cull(True)
    c3, w3 = h.scene_s3()   # extended program (N4 commands)
    ctx.cull_stats()
    assert np.array_equal(ctx.render(u, c3, w3, W, H), h.oracle_render(u, c3, w3, W, H)["rgba"])
    assert np.array_equal(ctx.render(u, 0, [], W, H), h.oracle_render(u, 0, [], W, H)["rgba"])   # empty program
    c1, w1 = h.scene_s1()   # 3 decoded instructions: the pre-pass would cost more than it saves (cfg2 1.07 -> 1.53 ms)
    assert np.array_equal(ctx.render(u, c1, w1, W, H), h.oracle_render(u, c1, w1, W, H)["rgba"])
    assert ctx.cull_stats()["culled_paints"] == 0
    with rm.Context(0, max_words=1024) as big:   # more than 64 decoded instructions
        big.set_mode(rm.MODE_EXACT)
        big.set_cull(True)
        cmd, words = h.scene_s2(n_leaves=80)
        assert np.array_equal(big.render(u, cmd, words, W, H), h.oracle_render(u, cmd, words, W, H)["rgba"])
        assert big.cull_stats()["culled_paints"] == 0


def test_cfg3_full_size_culled_identical_and_cheaper(rm, ctx):
    W, H = 3840, 2160
    cmd, words = h.scene_s2()
    u = h.orbit_uniforms(W, H, yaw=0.6, pitch=0.4)
    ctx.set_mode(rm.MODE_FAST)
    ctx.cull_stats()
    (a, ca), (b, cb) = on_off(ctx, u, cmd, words, W, H)
    st = ctx.cull_stats()
    assert np.array_equal(a, b) and np.array_equal(ca, cb)
    evals = int(cb[0]) + 4 * int(cb[1])
    print("cfg3 executed flops per evaluation:", st["executed_map_flops"] / evals, "of 798")
    assert st["executed_map_flops"] < 0.7 * 798 * evals   # measured: 509 of 798 flops per evaluation
