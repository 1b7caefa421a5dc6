"""N4 on the GPU: Plane and the Translation / Rotation / Scale Push / Pop commands through the
C ABI — the EXT instantiations of the production kernel (interpreted and run-time specialised) and of the verification
kernels.  Bars as for the reference instruction set: exact mode bit-identical to the CPU oracle (image, work
counters, per-sample AOVs, map_scene values), fast mode inside the north-star gates, bands and strips equal to the
full frame.  This file sorts after the other GPU test files on purpose: it covers the newest code.
"""
from pathlib import Path

import numpy as np
import pytest

import helpers as h
import rm_import
from test_gpu_parity import GATE_HITMASK, GATE_RGBA_FRAC, GATE_T_FRAC, gates

pytestmark = pytest.mark.gpu
ROOT = Path(__file__).resolve().parent.parent
GOLDEN = sorted((ROOT / "tests" / "golden" / "n4").glob("*.npz"))
E_UNSUPPORTED = -10


@pytest.fixture(scope="module")
def rm():
    return rm_import.load()


@pytest.fixture()
def ctx(rm):
    c = rm.Context(0)
    yield c
    c.close()


def load(p):
    z = np.load(p)
    return {k: z[k] for k in z.files}


@pytest.mark.parametrize("path", GOLDEN, ids=lambda p: p.stem)
@pytest.mark.parametrize("kernel", ["direct", "simple", "specialised"])
def test_exact_mode_reproduces_the_n4_golden_frames(rm, ctx, path, kernel):
    g = load(path)
    lim = g["limits"].tobytes()
    W, H, cmd, words, u = int(g["W"]), int(g["H"]), int(g["cmd_count"]), g["words"], g["uniforms"].tobytes()
    ctx.set_limits(*np.frombuffer(lim[:8], np.float32), int(np.frombuffer(lim[8:], np.uint32)[0]))
    ctx.set_mode(rm.MODE_EXACT)
    ctx.set_kernel(rm.KERNEL_SIMPLE if kernel == "simple" else rm.KERNEL_DIRECT)
    ctx.set_counters(True)
    ctx.prepare(u, cmd, words)
    if kernel == "specialised":
        ctx.specialize()
        assert ctx.spec_info()["active"]
    img = ctx.render(u, cmd, words, W, H)
    assert np.array_equal(img, g["rgba"]), gates(img, g["rgba"])
    assert np.array_equal(ctx.get_counters(), g["counters"])
    if kernel == "specialised":
        assert ctx.spec_info()["launches"] >= 1
    ctx.set_counters(False)
    dbg = ctx.paint_debug(W, H) if kernel == "simple" else ctx.paint_aov(W, H)   # AOVs of the kernel under test
    assert np.array_equal(dbg["cls"], g["cls"])
    assert np.array_equal(dbg["t"].view(np.uint32), g["t"].view(np.uint32))
    if kernel != "simple":
        assert bool(dbg["specialised"]) == (kernel == "specialised")


@pytest.mark.parametrize("seed", range(6))
def test_random_extended_trees_bit_identical(rm, ctx, seed):
    rng = np.random.default_rng(4321 + seed)
    u = h.orbit_uniforms(64, 48, yaw=0.5 * seed, pitch=0.3)
    n_ext = 0
    for n_leaves in (1, 2, 3, 6, 12):
        words, cmd = h.random_program(rng, n_leaves, ext=True)
        if len(words) > 255:
            continue
        n_ext += any(w in (2, 200, 202, 204) for w in words)
        pts = rng.uniform(-4, 4, (4096, 3)).astype(np.float32)
        ref_map = h.oracle_map(cmd, words, pts)
        ref = h.oracle_render(u, cmd, words, 64, 48)
        ctx.set_mode(rm.MODE_EXACT)
        ctx.set_kernel(rm.KERNEL_DIRECT)
        ctx.set_counters(True)
        ctx.prepare(u, cmd, words)
        got = ctx.map(pts)                                   # k_map (one point per thread)
        assert np.array_equal(got.view(np.uint32), ref_map.view(np.uint32)), (n_leaves, float(np.abs(got - ref_map).max()))
        img = ctx.render(u, cmd, words, 64, 48)              # production kernel, interpreter (4 or 8 rays per thread)
        assert np.array_equal(img, ref["rgba"]), (n_leaves, gates(img, ref["rgba"]))
        assert np.array_equal(ctx.get_counters(), ref["counters"])
        ctx.specialize()                                     # the same program compiled into the kernel
        img = ctx.render(u, cmd, words, 64, 48)
        assert ctx.spec_info()["active"]
        assert np.array_equal(img, ref["rgba"]), (n_leaves, gates(img, ref["rgba"]))
        assert np.array_equal(ctx.get_counters(), ref["counters"])
        ctx.set_kernel(rm.KERNEL_SIMPLE)
        assert np.array_equal(ctx.render(u, cmd, words, 64, 48), ref["rgba"])
        ctx.get_counters()
        ctx.set_mode(rm.MODE_FAST)
        got = ctx.map(pts)
        np.testing.assert_allclose(got, ref_map, rtol=0, atol=2e-5 * max(1.0, float(np.abs(ref_map).max())))
    assert n_ext >= 2


@pytest.mark.parametrize("specialise", [False, True])
def test_fast_mode_meets_the_north_star_gates_on_s3(rm, ctx, specialise):
    cmd, words = h.scene_s3()
    W, H = 384, 216
    u = h.orbit_uniforms(W, H, yaw=0.6, pitch=0.4)
    ref = h.oracle_render(u, cmd, words, W, H, aov=True)
    ctx.set_mode(rm.MODE_FAST)
    ctx.prepare(u, cmd, words)
    if specialise:
        ctx.specialize()
    img = ctx.render(u, cmd, words, W, H)
    aov = ctx.paint_aov(W, H)
    assert bool(aov["specialised"]) == specialise
    g = gates(img, ref["rgba"], aov["cls"], ref["cls"], aov["t"], ref["t"])
    print("S3", "specialised" if specialise else "interpreter", g)
    assert g["hitmask_pix"] <= GATE_HITMASK, g
    assert g["rgba_frac"] <= GATE_RGBA_FRAC, g
    assert g["t_frac"] <= GATE_T_FRAC, g


def test_short_extended_programs_take_the_four_ray_variant_and_long_ones_the_eight_ray_variant(rm, ctx):
    # <= 4 decoded instructions: 4 rays per thread; S3 (65 commands): 8.  Both against the oracle, ragged frame size.
    W, H = 83, 37
    u = h.orbit_uniforms(W, H, yaw=0.3, pitch=0.2)
    short = h.translated((0.5, 0.2, 0), (h.sphere((0, 0, 0), 1.0) + h.plane((0, 1, 0), 0.8) + [h.OP_SMOOTH_UNION, h.fbits(0.5)], 3))
    for words, cmd in (short, h.scene_s3()[::-1]):
        ref = h.oracle_render(u, cmd, words, W, H)
        ctx.set_mode(rm.MODE_EXACT)
        assert np.array_equal(ctx.render(u, cmd, words, W, H), ref["rgba"])


def test_bands_and_strips_equal_the_full_frame(rm, ctx):
    cmd, words = h.scene_s3()
    W, H = 96, 50
    u = h.orbit_uniforms(W, H, yaw=1.1, pitch=0.5)
    ctx.set_mode(rm.MODE_EXACT)
    full = ctx.render(u, cmd, words, W, H)
    assert np.array_equal(full, h.oracle_render(u, cmd, words, W, H)["rgba"])
    for row0, nrows in ((0, 7), (7, 20), (27, 23)):
        assert np.array_equal(ctx.render(u, cmd, words, W, H, row0=row0, nrows=nrows), full[row0:row0 + nrows])


def test_early_out_setting_is_ignored_for_extended_programs(rm, ctx):
    cmd, words = h.scene_s3()
    u = h.orbit_uniforms(64, 36, yaw=0.6, pitch=0.4)
    ctx.set_mode(rm.MODE_EXACT)
    ref = h.oracle_render(u, cmd, words, 64, 36)["rgba"]
    ctx.set_early_out(True)
    assert np.array_equal(ctx.render(u, cmd, words, 64, 36), ref)
    ctx.set_early_out(False)


@pytest.mark.parametrize("kernel", ["wavefront", "wavefront2"])
def test_earlier_generation_kernels_refuse_extended_programs(rm, ctx, kernel):
    cmd, words = h.scene_s3()
    u = h.orbit_uniforms(32, 32)
    ctx.set_kernel(rm.KERNEL_WAVEFRONT if kernel == "wavefront" else rm.KERNEL_WAVEFRONT2)
    ctx.prepare(u, cmd, words)
    with pytest.raises(rm.RmError) as e:
        ctx.paint(32, 32)
    assert e.value.code == E_UNSUPPORTED
    # the reference instruction set still runs on them, and the direct kernel takes the extended program
    c1, w1 = h.scene_s1()
    ctx.set_mode(rm.MODE_EXACT)
    assert np.array_equal(ctx.render(u, c1, w1, 32, 32), h.oracle_render(u, c1, w1, 32, 32)["rgba"])
    ctx.set_kernel(rm.KERNEL_DIRECT)
    assert np.array_equal(ctx.render(u, cmd, words, 32, 32), h.oracle_render(u, cmd, words, 32, 32)["rgba"])


def test_instruction_capacity_is_checked(rm):
    # a program that decodes to more instructions than it has words (tests/test_hostcheck_cpu.py shows it does): a context
    # sized for exactly its words refuses it, a larger one renders it — never a buffer overrun
    words, cmd = h.program_longer_than_its_words()
    u = h.orbit_uniforms(16, 16)
    with rm.Context(0, max_words=len(words)) as small:
        with pytest.raises(rm.RmError) as e:
            small.prepare(u, cmd, words)
        assert e.value.code == -2
    with rm.Context(0, max_words=2 * len(words)) as big:
        big.set_mode(rm.MODE_EXACT)
        assert np.array_equal(big.render(u, cmd, words, 16, 16), h.oracle_render(u, cmd, words, 16, 16)["rgba"])


def test_host_callback_renders_an_n4_tree(rm):
    import importlib
    host = importlib.import_module("ray_marching_b200.host")
    inner = host.binary(host.UNION, host.sphere((0, 0, 0), 0.5), host.box((0.6, 0, 0), (0.3, 0.3, 0.3)))
    R = h.rotation_matrix((0, 0, 1), 0.5)
    node = host.binary(host.SMOOTH_UNION, host.scale(host.rotation_matrix(host.translation(inner, (1, 0, 0)), R), 1.5),
                       host.plane((0, 1, 0), 1.2), 0.25)
    cmd, words = host.build_commands(node)
    o = host.Orbit((0, 0, 0), 5.0)
    o.set(0.6, 0.4)
    pos, q = o.camera()
    res = host.lib().rmh_resources_new(0)
    assert res >= 0
    img = host.callback_render(res, node, pos, q, 72, 40)
    ref = h.oracle_render(o.uniforms(72, 40), cmd, words, 72, 40)["rgba"]
    assert np.array_equal(img, ref)  # the host's contexts start in exact mode
    host.lib().rmh_reset()
