"""Oracle-vs-CUDA parity AT THE BASELINE SIZES: cfg2 on the whole 1920x1080 frame, cfg3 on 270 evenly
spaced rows of 3840x2160, cfg4 on 72 rows of 7680x4320 painted through rm_paint_strips_into, cfg5 on 4 of the 1024 orbit
poses at 1080p.  Exact mode must be byte-identical to the oracle (image, per-sample class, distance bit patterns); fast
mode must meet the three north-star gates — for the interpreter build and for the run-time specialised build of the
production kernel, from their own AOVs (rm_paint_aov)."""
import numpy as np
import pytest

import helpers as h
import rm_import

pytestmark = pytest.mark.gpu

GATE_HITMASK = 1e-4      # <= 0.01 % of pixels
GATE_RGBA_LSB = 2        # <= 2 LSB ...
GATE_RGBA_FRAC = 1e-3    # ... on >= 99.9 % of pixels
GATE_T_REL = 1e-4        # hit-distance relative error (on >= 99.9 % of the samples hit by both)
GATE_T_FRAC = 1e-3


@pytest.fixture(scope="module")
def rm():
    return rm_import.load()


@pytest.fixture()
def ctx(rm):
    c = rm.Context(0)
    yield c
    c.close()


def spaced_rows(H, n):
    return sorted({min(H - 1, int((i + 0.5) * H / n)) for i in range(n)})


def oracle_rows(u, cmd, words, W, H, rows):
    outs = [h.oracle_render(u, cmd, words, W, H, row0=r, nrows=1, aov=True) for r in rows]
    return (np.stack([o["rgba"][0] for o in outs]), np.stack([o["cls"][0] for o in outs]), np.stack([o["t"][0] for o in outs]))


def cuda_rows(ctx, W, H, rows):
    outs = [ctx.paint_aov(W, H, r, 1) for r in rows]
    return (np.stack([o["rgba"][0] for o in outs]), np.stack([o["cls"][0] for o in outs]), np.stack([o["t"][0] for o in outs]),
            all(o["specialised"] for o in outs))


def check_gates(img, cls, t, ref, what):
    rimg, rcls, rt = ref
    d = np.abs(img.astype(np.int32) - rimg.astype(np.int32)).max(-1)
    hm, hm_ref = cls == 2, rcls == 2
    both = hm & hm_ref
    rel = np.abs(t[both] - rt[both]) / np.maximum(np.abs(rt[both]), 1e-6)
    g = {"rgba_frac": float((d > GATE_RGBA_LSB).mean()), "rgba_max": int(d.max()), "hitmask_pix": float((hm != hm_ref).any(-1).mean()),
         "t_frac": float((rel > GATE_T_REL).mean()), "t_max": float(rel.max())}
    assert g["hitmask_pix"] <= GATE_HITMASK and g["rgba_frac"] <= GATE_RGBA_FRAC and g["t_frac"] <= GATE_T_FRAC, (what, g)
    return g


def check_identical(img, cls, t, ref, what):
    rimg, rcls, rt = ref
    assert np.array_equal(img, rimg), what
    assert np.array_equal(cls, rcls), what
    assert np.array_equal(t.view(np.uint32), rt.view(np.uint32)), what


def both_builds(rm, ctx):
    """(label, switch) for the interpreter build and the specialised build of the production kernel."""
    def interp():
        ctx.set_specialize(False)

    def spec():
        ctx.specialize()
    return (("interpreter", interp, False), ("specialised", spec, True))


def test_cfg2_whole_1920x1080_frame_against_the_oracle(rm, ctx):
    W, H = 1920, 1080
    cmd, words = h.scene_s1()
    u = h.orbit_uniforms(W, H, yaw=0.6, pitch=0.4)
    ref = h.oracle_render(u, cmd, words, W, H, aov=True)
    ref3 = (ref["rgba"], ref["cls"], ref["t"])
    for label, switch, is_spec in both_builds(rm, ctx):
        ctx.set_mode(rm.MODE_EXACT)
        ctx.prepare(u, cmd, words)
        switch()
        a = ctx.paint_aov(W, H)
        assert a["specialised"] == is_spec
        check_identical(a["rgba"], a["cls"], a["t"], ref3, f"cfg2 exact {label}")
        ctx.set_counters(True)
        img = ctx.render(u, cmd, words, W, H)          # the plain build (what is timed), with its work counters
        assert np.array_equal(img, ref["rgba"]) and np.array_equal(ctx.get_counters(), ref["counters"]), label
        ctx.set_counters(False)
        ctx.set_mode(rm.MODE_FAST)
        ctx.prepare(u, cmd, words)
        switch()
        a = ctx.paint_aov(W, H)
        assert a["specialised"] == is_spec
        check_gates(a["rgba"], a["cls"], a["t"], ref3, f"cfg2 fast {label}")
        img = ctx.render(u, cmd, words, W, H)
        assert np.array_equal(img, a["rgba"]), f"cfg2 fast {label}: plain build differs from its AOV build"


def test_cfg3_270_rows_of_3840x2160_against_the_oracle(rm, ctx):
    W, H = 3840, 2160
    cmd, words = h.scene_s2()
    u = h.orbit_uniforms(W, H, yaw=0.6, pitch=0.4)
    rows = spaced_rows(H, 270)
    ref = oracle_rows(u, cmd, words, W, H, rows)
    for label, switch, is_spec in both_builds(rm, ctx):
        ctx.set_mode(rm.MODE_EXACT)
        ctx.prepare(u, cmd, words)
        switch()
        img, cls, t, sp = cuda_rows(ctx, W, H, rows)
        assert sp == is_spec
        check_identical(img, cls, t, ref, f"cfg3 exact {label}")
        ctx.set_mode(rm.MODE_FAST)
        ctx.prepare(u, cmd, words)
        switch()
        img, cls, t, sp = cuda_rows(ctx, W, H, rows)
        assert sp == is_spec
        check_gates(img, cls, t, ref, f"cfg3 fast {label}")
        full = ctx.render(u, cmd, words, W, H)          # the timed build's whole frame: same rows, same bytes
        assert np.array_equal(full[rows], img), f"cfg3 fast {label}: plain build differs from its AOV build"


def test_cfg4_rows_of_7680x4320_through_strips_into_against_the_oracle(rm, ctx):
    import torch
    import importlib
    W, H, world = 7680, 4320, 8
    strip = importlib.import_module("ray_marching_b200.sharding").balanced_strip_rows(H, world)   # what bench.py runs: 12 rows
    assert strip == 12
    cmd, words = h.scene_s2()
    u = h.orbit_uniforms(W, H, yaw=0.6, pitch=0.4)
    rows = spaced_rows(H, 72)
    ref_img = np.stack([h.oracle_render(u, cmd, words, W, H, row0=r, nrows=1)["rgba"][0] for r in rows])
    frame = torch.zeros((H, W, 4), dtype=torch.uint8, device="cuda:0")
    for mode, spec in ((rm.MODE_EXACT, False), (rm.MODE_FAST, False), (rm.MODE_FAST, True)):
        frame.zero_()
        ctx.set_mode(mode)
        ctx.prepare(u, cmd, words)
        if spec:
            ctx.specialize()
        else:
            ctx.set_specialize(False)
        for r in range(world):      # the 8 shards of the config, one after the other on this GPU, each into the whole frame
            ctx.paint_strips_into(W, H, strip, r, world, frame.data_ptr(), W * 4)
        ctx.sync()
        got = frame[torch.as_tensor(rows, device="cuda:0")].cpu().numpy()
        if mode == rm.MODE_EXACT:
            assert np.array_equal(got, ref_img)
        else:
            d = np.abs(got.astype(np.int32) - ref_img.astype(np.int32)).max(-1)
            assert float((d > GATE_RGBA_LSB).mean()) <= GATE_RGBA_FRAC, (spec, int(d.max()), float((d > GATE_RGBA_LSB).mean()))
        assert int(frame[..., 3].min()) == 255      # every row of the frame was stored by some shard


def test_cfg5_four_of_the_1024_orbit_poses_at_1080p_against_the_oracle(rm, ctx):
    W, H = 1920, 1080
    cmd, words = h.scene_s1()
    ks = (0, 257, 514, 1023)
    us = [h.orbit_uniforms(W, H, yaw=2 * np.pi * k / 1024, pitch=0.4) for k in ks]
    refs = [h.oracle_render(u, cmd, words, W, H, aov=True) for u in us]
    buf = rm.PinnedBuffer(len(ks) * W * H * 4)
    for mode in (rm.MODE_EXACT, rm.MODE_FAST):
        for spec in (False, True):
            ctx.set_mode(mode)
            ctx.prepare(us[0], cmd, words)
            if spec:
                ctx.specialize()
            else:
                ctx.set_specialize(False)
            ctx.paint_poses_host(b"".join(us), len(ks), W, H, buf.ptr)     # the pose-batch path of the config
            got = buf.array.reshape(len(ks), H, W, 4)
            for i, ref in enumerate(refs):
                if mode == rm.MODE_EXACT:
                    assert np.array_equal(got[i], ref["rgba"]), (spec, ks[i])
                else:
                    ctx.prepare_uniforms(us[i])
                    a = ctx.paint_aov(W, H)
                    assert a["specialised"] == spec and np.array_equal(a["rgba"], got[i])
                    check_gates(a["rgba"], a["cls"], a["t"], (ref["rgba"], ref["cls"], ref["t"]), f"cfg5 pose {ks[i]} spec={spec}")
    buf.free()
