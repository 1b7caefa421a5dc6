"""The opt-in early-out of the interpreter (rm_set_early_out, csrc/rm_wavefront_common.cuh) skips the rest of a
primitive+operator instruction when a lower bound of the primitive proves the accumulator unchanged.  The claim is
bit-identity: image and work counters with it ON must equal those with it OFF in both arithmetic modes, and exact mode
must still equal the CPU oracle."""
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
    c.set_kernel(rm.KERNEL_DIRECT)
    yield c
    c.close()


def on_off(ctx, u, cmd, words, W, H):
    out = []
    for eo in (False, True):
        ctx.set_early_out(eo)
        ctx.set_counters(True)
        img = ctx.render(u, cmd, words, W, H)
        out.append((img, ctx.get_counters()))
        ctx.set_counters(False)
    ctx.set_early_out(False)
    return out


@pytest.mark.parametrize("name,W,H,yaw,pitch,radius", [("S0", 320, 240, 0.0, 0.0, 5.0), ("S1", 480, 270, 0.6, 0.4, 5.0),
                                                       ("S2", 384, 216, 0.6, 0.4, 5.0), ("S2", 300, 170, 2.4, 0.1, 3.0),
                                                       ("S2", 256, 144, 1.1, 0.7, 1.2)])
@pytest.mark.parametrize("mode", ["exact", "fast"])
def test_on_equals_off_bit_for_bit(rm, ctx, name, W, H, yaw, pitch, radius, mode):
    cmd, words = h.SCENES[name]()
    u = h.orbit_uniforms(W, H, yaw=yaw, pitch=pitch, radius=radius)
    ctx.set_mode(rm.MODE_EXACT if mode == "exact" else rm.MODE_FAST)
    (a, ca), (b, cb) = on_off(ctx, u, cmd, words, W, H)
    assert np.array_equal(a, b), int((a != b).any(-1).sum())
    assert np.array_equal(ca, cb), (ca, cb)


@pytest.mark.parametrize("seed", range(5))
def test_random_trees_on_equals_oracle(rm, ctx, seed):
    """random left/right-leaning trees with unions, subtractions, intersections and smooth unions of several k"""
    rng = np.random.default_rng(777 + seed)
    u = h.orbit_uniforms(72, 56, yaw=0.5 * seed, pitch=0.3)
    ctx.set_mode(rm.MODE_EXACT)
    ctx.set_early_out(True)
    for n_leaves in (2, 5, 11, 24):
        words, cmd = random_program(rng, n_leaves, right_bias=0.25)
        if len(words) > 255:
            continue
        img = ctx.render(u, cmd, words, 72, 56)
        assert np.array_equal(img, h.oracle_render(u, cmd, words, 72, 56)["rgba"]), n_leaves
    ctx.set_early_out(False)


def test_limits_and_full_size(rm, ctx):
    cmd, words = h.scene_s2()
    u = h.orbit_uniforms(96, 64, yaw=0.6, pitch=0.4)
    ctx.set_mode(rm.MODE_EXACT)
    for lim in [(0.05, 50.0, 20), (0.001, 100.0, 7), (0.3, 100.0, 100), (0.01, 6.0, 100)]:
        ctx.set_limits(*lim)
        (a, ca), (b, cb) = on_off(ctx, u, cmd, words, 96, 64)
        assert np.array_equal(a, b) and np.array_equal(ca, cb), lim
        assert np.array_equal(b, h.oracle_render(u, cmd, words, 96, 64, limits=h.limits_bytes(*lim))["rgba"]), lim
    ctx.set_limits(0.01, 100.0, 100)
    W, H = 3840, 2160
    u = h.orbit_uniforms(W, H, yaw=0.6, pitch=0.4)
    ctx.set_mode(rm.MODE_FAST)
    (a, ca), (b, cb) = on_off(ctx, u, cmd, words, W, H)
    assert np.array_equal(a, b) and np.array_equal(ca, cb)
