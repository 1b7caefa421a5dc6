"""Program specialisation on the GPU (rm_specialize / rm_set_specialize): map_scene of the prepared command stream is
compiled at run time into straight-line code around the production kernel body.
Bars: exact mode bit-identical to the CPU oracle (image and work counters) — the arithmetic per ray is the
interpreter's, operation for operation; fast mode inside the north-star gates against the oracle and within 1 LSB of the
interpreter; every band / strip / pose-batch path unchanged; falls back to the interpreter when the words change, when
NVRTC is missing, and for programs too long for the instruction cache."""
import os
import subprocess
import sys
from pathlib import Path

import numpy as np
import pytest

import helpers as h
import rm_import
from test_gpu_parity import GATE_HITMASK, GATE_RGBA_FRAC, GATE_RGBA_LSB, cases, gates, random_program

pytestmark = pytest.mark.gpu
ROOT = Path(__file__).resolve().parent.parent


@pytest.fixture(scope="module")
def rm():
    return rm_import.load()


@pytest.fixture()
def ctx(rm):
    c = rm.Context(0)
    yield c
    c.close()


@pytest.mark.parametrize("name,W,H,yaw,pitch", cases())
def test_specialised_exact_mode_is_bit_identical_to_the_oracle(rm, ctx, name, W, H, yaw, pitch):
    cmd, words = h.SCENES[name]()
    u = h.orbit_uniforms(W, H, yaw=yaw, pitch=pitch)
    ref = h.oracle_render(u, cmd, words, W, H, aov=True)
    ctx.set_mode(rm.MODE_EXACT)
    ctx.set_counters(True)
    ctx.prepare(u, cmd, words)
    ctx.specialize()
    info = ctx.spec_info()
    assert info["active"] and info["rays_per_thread"] == (2 if name == "S2" else 4) and info["compile_ms"] > 0
    n0 = info["launches"]
    img = ctx.render(u, cmd, words, W, H)
    cnt = ctx.get_counters()
    assert ctx.spec_info()["launches"] == n0 + 1  # the specialised kernel is what ran
    assert np.array_equal(img, ref["rgba"]), gates(img, ref["rgba"])
    assert np.array_equal(cnt, ref["counters"]), (cnt, ref["counters"])


@pytest.mark.parametrize("name,W,H,yaw,pitch", cases())
def test_specialised_fast_mode_meets_the_gates_and_tracks_the_interpreter(rm, ctx, name, W, H, yaw, pitch):
    cmd, words = h.SCENES[name]()
    u = h.orbit_uniforms(W, H, yaw=yaw, pitch=pitch)
    ref = h.oracle_render(u, cmd, words, W, H)
    ctx.set_mode(rm.MODE_FAST)
    interp = ctx.render(u, cmd, words, W, H)
    ctx.specialize()
    img = ctx.render(u, cmd, words, W, H)
    assert ctx.spec_info()["active"] and ctx.spec_info()["launches"] == 1
    g = gates(img, ref["rgba"])
    assert g["rgba_frac"] <= GATE_RGBA_FRAC, g
    d = np.abs(img.astype(int) - interp.astype(int)).max(-1)
    assert d.max() <= 1 and (d > 0).mean() <= 1e-3, (int(d.max()), float((d > 0).mean()))
    # hit mask: the class of every sample comes out of the same comparisons; count disagreeing PIXELS through the image
    # (a sample that flips between hit and miss changes its pixel by far more than 2 LSB only at silhouettes)
    assert (d > GATE_RGBA_LSB).mean() <= GATE_HITMASK


@pytest.mark.parametrize("seed", range(4))
def test_specialised_random_trees_exact(rm, ctx, seed):
    """All opcodes, mixed tree shapes (operands that are subtrees, stack depth > 2), several k values."""
    rng = np.random.default_rng(4321 + seed)
    W, H = 96, 64
    u = h.orbit_uniforms(W, H, yaw=0.3 + seed, pitch=0.35)
    ctx.set_mode(rm.MODE_EXACT)
    ctx.set_counters(True)
    for n_leaves in (1, 2, 5, 12, 24):
        words, cmd = random_program(rng, n_leaves, right_bias=0.5)
        if len(words) > 255:
            continue
        ref = h.oracle_render(u, cmd, words, W, H)
        ctx.prepare(u, cmd, words)
        ctx.specialize()
        assert ctx.spec_info()["active"]
        ctx.paint(W, H)
        img = np.empty((H, W, 4), np.uint8)
        ctx.download(img)
        assert np.array_equal(img, ref["rgba"]), (n_leaves, gates(img, ref["rgba"]))
        assert np.array_equal(ctx.get_counters(), ref["counters"])


def test_empty_program_specialises_to_max_dist(rm, ctx):
    u = h.orbit_uniforms(64, 48, pitch=0.3)
    ctx.set_mode(rm.MODE_EXACT)
    ctx.prepare(u, 0, [])
    ctx.specialize()
    img = ctx.render(u, 0, [], 64, 48)
    assert ctx.spec_info()["launches"] == 1
    assert np.array_equal(img, h.oracle_render(u, 0, [], 64, 48)["rgba"])


def test_bands_strips_and_pose_batches_run_specialised(rm, ctx):
    W, H = 200, 77  # ragged: not a multiple of any tile
    cmd, words = h.scene_s2()
    u = h.orbit_uniforms(W, H, yaw=0.6, pitch=0.4)
    ctx.set_mode(rm.MODE_FAST)
    ctx.prepare(u, cmd, words)
    ctx.specialize()
    full = ctx.render(u, cmd, words, W, H)
    assert ctx.spec_info()["rays_per_thread"] == 2
    # row bands
    for row0, n in ((0, 1), (5, 31), (70, 7), (0, 77)):
        band = ctx.render(u, cmd, words, W, H, row0, n)
        assert np.array_equal(band, full[row0:row0 + n])
    # interleaved strips stored at their frame rows
    import torch
    frame = torch.zeros((H, W, 4), dtype=torch.uint8, device="cuda:0")
    for first in range(3):
        ctx.paint_strips_into(W, H, 4, first, 3, frame.data_ptr())
    ctx.sync()
    assert np.array_equal(frame.cpu().numpy(), full)
    # pose batch returned to host
    us = [h.orbit_uniforms(W, H, yaw=0.9 * k, pitch=0.4) for k in range(5)]
    singles = [ctx.render(p, cmd, words, W, H) for p in us]
    buf = rm.PinnedBuffer(5 * W * H * 4)
    n0 = ctx.spec_info()["launches"]
    ctx.paint_poses_host(b"".join(us), 5, W, H, buf.ptr)
    assert ctx.spec_info()["launches"] == n0 + 5
    for k in range(5):
        assert np.array_equal(buf.array.reshape(5, H, W, 4)[k], singles[k])
    buf.free()


def test_new_words_go_back_to_the_interpreter_and_the_policy_recompiles(rm, ctx):
    W, H = 96, 64
    u = h.orbit_uniforms(W, H, yaw=0.3, pitch=0.2)
    ctx.set_mode(rm.MODE_EXACT)
    cmd, words = h.scene_s1()
    ctx.prepare(u, cmd, words)
    ctx.specialize()
    ctx.render(u, cmd, words, W, H)
    assert ctx.spec_info()["launches"] == 1
    cmd2, words2 = h.scene_s0()
    img = ctx.render(u, cmd2, words2, W, H)       # different words: interpreter, no compile
    info = ctx.spec_info()
    assert not info["active"] and info["launches"] == 1
    assert np.array_equal(img, h.oracle_render(u, cmd2, words2, W, H)["rgba"])
    ctx.set_specialize(True)                       # policy: every program, compiled at first paint
    img = ctx.render(u, cmd2, words2, W, H)
    info = ctx.spec_info()
    assert info["active"] and info["launches"] == 2
    assert np.array_equal(img, h.oracle_render(u, cmd2, words2, W, H)["rgba"])
    compiles = info["compiles"]
    ctx.render(u, cmd, words, W, H)                # S1 again: already in the process-wide cache
    assert ctx.spec_info()["compiles"] == compiles and ctx.spec_info()["launches"] == 3
    # a second context shares the cache
    with rm.Context(0) as other:
        other.set_mode(rm.MODE_EXACT)
        other.prepare(u, cmd, words)
        other.specialize()
        assert other.spec_info()["compiles"] == 0 and other.spec_info()["active"]


def test_too_long_a_program_stays_on_the_interpreter(rm):
    n = 120
    words = h.box((0, 0, 0), (1, 1, 1))
    for i in range(n - 1):
        words = words + h.box((0.02 * i, 0, 0), (0.5, 0.5, 0.5)) + [103, h.fbits(0.25)]
    cmd = 2 * n - 1
    W, H = 48, 32
    u = h.orbit_uniforms(W, H, yaw=0.3, pitch=0.2)
    with rm.Context(0, max_words=2048) as ctx:
        ctx.set_mode(rm.MODE_EXACT)
        ctx.prepare(u, cmd, words)
        with pytest.raises(rm.RmError) as e:
            ctx.specialize()
        assert e.value.code == -9 and "too long" in str(e.value)
        img = ctx.render(u, cmd, words, W, H)
        assert ctx.spec_info()["launches"] == 0
        ctx.set_specialize(True)  # the policy fails silently and says why
        img2 = ctx.render(u, cmd, words, W, H)
        assert np.array_equal(img, img2) and "too long" in ctx.spec_info()["status"]


def test_without_nvrtc_specialize_reports_and_painting_continues():
    """RM_NVRTC_PATH names an exact file; a missing one must give RM_E_JIT, not a crash and not a silent CPU path."""
    code = (
        "import sys; sys.path.insert(0, %r); sys.path.insert(0, %r)\n"
        "import numpy as np, helpers as h, rm_import\n"
        "rm = rm_import.load()\n"
        "ctx = rm.Context(0); ctx.set_mode(rm.MODE_EXACT)\n"
        "cmd, words = h.scene_s1(); u = h.orbit_uniforms(64, 48)\n"
        "ctx.prepare(u, cmd, words)\n"
        "try:\n"
        "    ctx.specialize(); print('NOERROR')\n"
        "except rm.RmError as e:\n"
        "    print('CODE', e.code, 'NVRTC is not available' in str(e))\n"
        "img = ctx.render(u, cmd, words, 64, 48)\n"
        "print('SAME', bool(np.array_equal(img, h.oracle_render(u, cmd, words, 64, 48)['rgba'])), ctx.spec_info()['launches'])\n"
    ) % (str(ROOT), str(ROOT / "tests"))
    env = dict(os.environ, RM_NVRTC_PATH="/nonexistent/libnvrtc.so")
    r = subprocess.run([sys.executable, "-c", code], env=env, capture_output=True, text=True, timeout=300)
    assert "CODE -9 True" in r.stdout, r.stdout + r.stderr
    assert "SAME True 0" in r.stdout, r.stdout + r.stderr


def test_group_specialises_once_for_all_members(rm):
    import torch
    have = max(1, torch.cuda.device_count())
    W, H = 256, 150
    cmd, words = h.scene_s2()
    u = h.orbit_uniforms(W, H, yaw=1.1, pitch=0.4)
    with rm.Context(0) as ctx:
        ctx.set_mode(rm.MODE_EXACT)
        full = ctx.render(u, cmd, words, W, H)
    with rm.Group([k % have for k in range(3)]) as g:
        g.set_mode(rm.MODE_EXACT)
        g.prepare(u, cmd, words)
        g.specialize()
        got = np.zeros((H, W, 4), np.uint8)
        g.paint_bands(W, H, 16, got)
    assert np.array_equal(got, full)


@pytest.mark.parametrize("name,W,H,yaw,pitch", cases())
@pytest.mark.parametrize("specialised", [False, True])
def test_production_kernel_own_aovs_meet_all_three_gates(rm, ctx, name, W, H, yaw, pitch, specialised):
    """rm_paint_aov: classes and hit distances written by the production kernel's OWN march (interpreter build and
    run-time specialised build), not by the verification kernel.  Exact mode: bit-identical to the oracle.  Fast mode:
    hit-mask, RGBA8 and hit-distance gates of the north star."""
    from test_gpu_parity import GATE_T_FRAC
    cmd, words = h.SCENES[name]()
    u = h.orbit_uniforms(W, H, yaw=yaw, pitch=pitch)
    ref = h.oracle_render(u, cmd, words, W, H, aov=True)
    for mode in (rm.MODE_EXACT, rm.MODE_FAST):
        ctx.set_mode(mode)
        ctx.prepare(u, cmd, words)
        if specialised:
            ctx.specialize()
        else:
            ctx.set_specialize(False)
        out = ctx.paint_aov(W, H)
        assert out["specialised"] == specialised
        assert np.array_equal(out["rgba"], ctx.render(u, cmd, words, W, H))  # the AOV build paints the same image
        if mode == rm.MODE_EXACT:
            assert np.array_equal(out["cls"], ref["cls"])
            assert np.array_equal(out["t"].view(np.uint32), ref["t"].view(np.uint32))
            assert np.array_equal(out["rgba"], ref["rgba"])
        else:
            g = gates(out["rgba"], ref["rgba"], out["cls"], ref["cls"], out["t"], ref["t"])
            print(name, "specialised" if specialised else "interpreter", g)
            assert g["hitmask_pix"] <= GATE_HITMASK, g
            assert g["rgba_frac"] <= GATE_RGBA_FRAC, g
            assert g["t_frac"] <= GATE_T_FRAC, g


def test_aov_bands_and_ragged_sizes(rm, ctx):
    W, H = 77, 45
    cmd, words = h.scene_s2()
    u = h.orbit_uniforms(W, H, yaw=0.6, pitch=0.4)
    ref = h.oracle_render(u, cmd, words, W, H, aov=True)
    ctx.set_mode(rm.MODE_EXACT)
    ctx.prepare(u, cmd, words)
    ctx.specialize()
    for row0, n in ((0, 45), (7, 13), (44, 1)):
        out = ctx.paint_aov(W, H, row0, n)
        assert out["specialised"]
        assert np.array_equal(out["cls"], ref["cls"][row0:row0 + n])
        assert np.array_equal(out["t"].view(np.uint32), ref["t"][row0:row0 + n].view(np.uint32))
        assert np.array_equal(out["rgba"], ref["rgba"][row0:row0 + n])


def test_background_policy_never_stalls_and_switches_over(rm, ctx):
    """rm_set_specialize(2): the first paints of a new program go through the interpreter while a worker thread compiles;
    a later paint adopts the kernel.  Exact mode, so every frame on either path must equal the oracle."""
    import time
    W, H = 96, 64
    u = h.orbit_uniforms(W, H, yaw=0.3, pitch=0.2)
    # a program STRUCTURE nobody has compiled in this process yet (the cache is process-wide and keyed by structure: the
    # zero pattern of the parameters is part of it)
    words = (h.sphere((0.123, 0.0, -0.1), 0.9) + h.box((0.4, 0.1, 0.0), (0.5, 0.6, 0.7)) + [103, h.fbits(0.37)]
             + h.sphere((0.0, 0.0, 0.3), 0.4) + [102] + h.box((0.0, 0.2, 0.0), (0.3, 0.3, 0.9)) + [101])
    ncmd = 7
    ref = h.oracle_render(u, ncmd, words, W, H)["rgba"]
    ctx.set_mode(rm.MODE_EXACT)
    ctx.set_specialize(2)
    t0 = time.perf_counter()
    img = ctx.render(u, ncmd, words, W, H)
    first_ms = (time.perf_counter() - t0) * 1e3
    assert np.array_equal(img, ref)
    assert ctx.spec_info()["launches"] == 0           # painted by the interpreter, the compile had only just started
    assert first_ms < 100, first_ms                   # no compile in the paint path (a compile is 150-600 ms)
    deadline = time.time() + 30
    while ctx.spec_info()["launches"] == 0 and time.time() < deadline:
        assert np.array_equal(ctx.render(u, ncmd, words, W, H), ref)
        time.sleep(0.02)
    info = ctx.spec_info()
    assert info["active"] and info["launches"] >= 1 and info["compiles"] == 1 and ctx.spec_tier == 1
    # other VALUES of the same structure: adopted from the cache at once, no compile, no interpreter frame
    w2 = list(words)
    w2[4] = h.fbits(0.55)
    n0 = info["launches"]
    assert np.array_equal(ctx.render(u, ncmd, w2, W, H), h.oracle_render(u, ncmd, w2, W, H)["rgba"])
    assert ctx.spec_info()["launches"] == n0 + 1 and ctx.spec_info()["compiles"] == 1
    # words that change while a compile is in flight: its result is dropped, nothing crashes, frames stay right
    for i in range(4):
        w2 = h.sphere((0.01 * i, 0.2, -0.1), 0.8 + 0.01 * i)
        assert np.array_equal(ctx.render(u, 1, w2, W, H), h.oracle_render(u, 1, w2, W, H)["rgba"])
    with pytest.raises(rm.RmError):
        ctx.set_specialize(3)



# ---- round 2: kernels are compiled per program STRUCTURE; values travel as kernel arguments ------------
def test_dragging_a_radius_for_1000_frames_compiles_once(rm, ctx):
    """The reference re-flattens and re-uploads the program every frame (renderer.rs:224-239) and an editor drags values:
    1000 frames, one radius changing per frame -> ONE compile, every frame specialised and byte-identical to the oracle."""
    W, H = 48, 27
    u = h.orbit_uniforms(W, H, yaw=0.6, pitch=0.4)
    ctx.set_mode(rm.MODE_EXACT)
    ctx.set_specialize(1)
    c0 = None
    for f in range(1000):
        r = 0.4 + 0.0005 * f
        words = h.box((0.05, 0.0, 0.0), (1, 1, 1)) + h.sphere((0.0, 1.0, 0.0), 1.0) + [h.OP_UNION] + h.sphere((0.6, 0.3, 1.0), r) + [h.OP_SUBTRACTION]
        img = ctx.render(u, 5, words, W, H)
        info = ctx.spec_info()
        if c0 is None:
            c0 = info["compiles"]            # 1, or 0 if an earlier test compiled this structure
            assert c0 <= 1
        assert info["active"] and info["launches"] == f + 1 and info["compiles"] == c0, (f, info)
        assert ctx.spec_tier == 1            # the words never stay put: no literal-tier compile either
        if f % 7 == 0 or f == 999:
            assert np.array_equal(img, h.oracle_render(u, 5, words, W, H)["rgba"]), f
    assert ctx.program_uploads >= 1000


def test_streaming_more_structures_than_the_cache_holds_keeps_specialising(rm, ctx):
    """LRU eviction instead of a permanent cap: with room for 6 kernels, 20 distinct structures are all painted
    specialised and right; evicted kernels are unloaded (cudaLibraryUnload) once no context paints with them."""
    W, H = 64, 36
    u = h.orbit_uniforms(W, H, yaw=0.6, pitch=0.4)
    ctx.set_mode(rm.MODE_EXACT)
    ctx.set_specialize(1)
    ctx.set_spec_literal_after(0)
    before = rm.spec_cache()
    try:
        rm.spec_cache(6)
        ev0 = rm.spec_cache()["evictions"]
        rng = np.random.default_rng(77)
        seen = []
        for i in range(20):
            words, cmd = h.random_program(rng, 2 + i % 5, ext=(i % 3 == 0))
            words = list(words)
            for j in range(i + 1):   # i + 1 more leaves: every structure of the stream has its own length
                words += h.sphere((0.1 * i, 0.05 * j, 0), 0.3) + [h.OP_UNION]
            cmd += 2 * (i + 1)
            img = ctx.render(u, cmd, words, W, H)
            info = ctx.spec_info()
            assert info["active"], (i, info["status"])
            assert np.array_equal(img, h.oracle_render(u, cmd, words, W, H)["rgba"]), i
            seen.append((cmd, words))
        st = rm.spec_cache()
        assert st["entries"] <= 6 and st["evictions"] - ev0 >= 10, st
        # an evicted structure comes back with a fresh compile and still paints right
        cmd, words = seen[0]
        n0 = ctx.spec_info()["compiles"]
        assert np.array_equal(ctx.render(u, cmd, words, W, H), h.oracle_render(u, cmd, words, W, H)["rgba"])
        assert ctx.spec_info()["compiles"] == n0 + 1 and ctx.spec_info()["active"]
    finally:
        rm.spec_cache(before["capacity"])


def test_words_that_stay_put_are_promoted_to_the_literal_tier_in_the_background(rm, ctx):
    import time
    W, H = 96, 54
    u = h.orbit_uniforms(W, H, yaw=0.6, pitch=0.4)
    words = h.box((0.07, 0.0, 0.0), (0.9, 1, 1)) + h.sphere((0.0, 1.0, 0.0), 1.0) + [h.OP_SMOOTH_UNION, h.fbits(0.31)]
    ref = h.oracle_render(u, 3, words, W, H)["rgba"]
    ctx.set_mode(rm.MODE_EXACT)
    ctx.set_specialize(1)
    ctx.set_spec_literal_after(4)
    tiers = []
    deadline = time.time() + 60
    while time.time() < deadline:
        t0 = time.perf_counter()
        img = ctx.render(u, 3, words, W, H)
        ms = (time.perf_counter() - t0) * 1e3
        assert np.array_equal(img, ref)
        tiers.append(ctx.spec_tier)
        if len(tiers) > 1:
            assert ms < 100, (len(tiers), ms)       # only the very first paint may wait for a compile (policy 1)
        if tiers[-1] == 2:
            break
        time.sleep(0.01)
    assert tiers[0] == 1 and tiers[-1] == 2 and tiers[:4] == [1, 1, 1, 1], tiers
    assert np.array_equal(ctx.render(u, 3, words, W, H), ref)     # painted by the literal kernel
    c0 = ctx.spec_info()["compiles"]
    w2 = list(words)
    w2[4] = h.fbits(0.8)                                           # a value changes: structure kernel again, no compile
    assert np.array_equal(ctx.render(u, 3, w2, W, H), h.oracle_render(u, 3, w2, W, H)["rgba"])
    assert ctx.spec_tier == 1 and ctx.spec_info()["compiles"] == c0
    ctx.set_spec_literal_after(0)                                   # structure tier only
    for _ in range(6):
        assert np.array_equal(ctx.render(u, 3, words, W, H), ref)
    assert ctx.spec_tier == 1


def test_specialize_compiles_both_tiers_at_once_and_fast_mode_tiers_agree(rm, ctx):
    W, H = 384, 216
    cmd, words = h.scene_s2()
    u = h.orbit_uniforms(W, H, yaw=0.6, pitch=0.4)
    ctx.set_mode(rm.MODE_FAST)
    ctx.prepare(u, cmd, words)
    ctx.specialize()
    assert ctx.spec_tier == 2
    lit = ctx.render(u, cmd, words, W, H)
    ctx.set_spec_literal_after(0)
    assert ctx.spec_tier == 1
    par = ctx.render(u, cmd, words, W, H)
    ctx.set_specialize(False)
    interp = ctx.render(u, cmd, words, W, H)
    for a, b in ((lit, par), (par, interp)):
        d = np.abs(a.astype(np.int16) - b.astype(np.int16)).max(-1)
        assert d.max() <= 1 and (d > 0).mean() <= 2e-3, (int(d.max()), float((d > 0).mean()))


def _s2_variant(scale):
    """S2's structure with other values: every radius scaled (no zero appears or disappears, so the structure is the same)."""
    cmd, words = h.scene_s2()
    w = np.array(words, np.uint32)
    f = w.view(np.float32)
    i = 0
    while i < len(w):
        op = int(w[i])
        if op == 0:
            f[i + 4] *= scale
            i += 5
        elif op == 1:
            f[i + 4:i + 7] *= scale
            i += 7
        elif op == h.OP_SMOOTH_UNION:
            i += 2
        else:
            i += 1
    return cmd, [int(x) for x in w]


def test_deep_structure_tier_reads_its_values_from_the_modules_constant_block(rm):
    """The structure tier of a deep program keeps its parameters in a __constant__ block of the generated module, ONE per
    device, shared by every context that paints that structure (rm_jit.cu launch_spec).  Two contexts of one GPU painting
    the same structure with DIFFERENT values, launches interleaved on their own streams with no synchronisation in between,
    and one context alternating values frame by frame, must each get their own frame — bit-identical to the oracle."""
    W, H = 512, 288
    u = h.orbit_uniforms(W, H, yaw=0.6, pitch=0.4)
    progs = [_s2_variant(s) for s in (1.0, 0.8, 1.15)]
    assert rm.spec_source(*progs[0]) == rm.spec_source(*progs[1]) == rm.spec_source(*progs[2])
    assert "__constant__ rm::SpecParams g_rm_spec_P;" in rm.spec_source(*progs[0])
    rows = [0, 57, 143, 144, 287]
    refs = [np.stack([h.oracle_render(u, c, w, W, H, row0=r, nrows=1)["rgba"][0] for r in rows]) for c, w in progs]
    with rm.Context(0) as a, rm.Context(0) as b:
        for c in (a, b):
            c.set_mode(rm.MODE_EXACT)
            c.set_specialize(1)
            c.set_spec_literal_after(0)          # never promoted: every paint below runs the structure kernel
        out_a, out_b = np.zeros((H, W, 4), np.uint8), np.zeros((H, W, 4), np.uint8)
        for it in range(6):
            ia, ib = (0, 2)[it % 2], 1           # a alternates between two value sets, b keeps its own
            a.prepare(u, *progs[ia])
            a.paint(W, H)                        # asynchronous, a's stream
            b.prepare(u, *progs[ib])
            b.paint(W, H)                        # asynchronous, b's stream: a's kernel may still be running
            assert a.spec_tier == 1 and b.spec_tier == 1 and a.spec_info()["rays_per_thread"] == 2
            if it >= 4:                          # the last two rounds are read back (a: one frame of each value set)
                a.download(out_a)
                b.download(out_b)
                assert np.array_equal(out_a[rows], refs[ia]), (it, "a")
                assert np.array_equal(out_b[rows], refs[ib]), (it, "b")
        assert a.spec_info()["compiles"] + b.spec_info()["compiles"] <= 1   # one structure: at most one compile (0 if cached)
        # a pose batch (two streams inside the library) through the same kernel, in fast mode against single paints
        a.set_mode(rm.MODE_FAST)
        poses = [h.orbit_uniforms(W, H, yaw=0.4 * k, pitch=0.3) for k in range(4)]
        buf = rm.PinnedBuffer(len(poses) * W * H * 4)
        a.prepare(poses[0], *progs[2])
        a.specialize()
        a.set_spec_literal_after(0)
        a.paint_poses_host(b"".join(poses), len(poses), W, H, buf.ptr)
        got = buf.array.reshape(len(poses), H, W, 4).copy()
        buf.free()
        assert a.spec_tier == 1
        for k, pu in enumerate(poses):
            b.set_mode(rm.MODE_FAST)
            b.set_specialize(False)
            ref = b.render(pu, *progs[2], W, H)  # the interpreter, another context
            d = np.abs(got[k].astype(np.int32) - ref.astype(np.int32)).max(-1)
            assert float((d > 1).mean()) <= 4e-3, (k, int(d.max()))


@pytest.mark.parametrize("k", [0.05, 1.0, 3.99, 4.0, 4.01, 9.0])
def test_fast_smooth_union_as_a_saturate_up_to_k_4_and_the_general_form_beyond(rm, ctx, k):
    """Fast mode clamps the smooth union's blend term with the FFMA's .SAT modifier when k*s <= 1 (k <= 4) — the same value
    as max(.,0) there — and keeps max(.,0) beyond.  A 10-leaf chain (8-ray interpreter with the common tail, both compiled
    tiers) on either side of the switch: every kernel meets the image gate against the oracle (the one-sample kernel never
    uses the saturate), and the production kernel's three builds differ only by multiply-add contraction."""
    W, H = 192, 108
    u = h.orbit_uniforms(W, H, yaw=0.5, pitch=0.35, radius=7.0)
    words = h.sphere((-2.0, 0.0, 0.0), 0.7)
    for i in range(1, 10):
        leaf = h.box((-2.0 + 0.45 * i, 0.15 * (i % 3), 0.1 * i), (0.35, 0.3, 0.25)) if i % 2 else h.sphere((-2.0 + 0.45 * i, 0.2, 0.0), 0.45)
        words = words + leaf + [h.OP_SMOOTH_UNION, h.fbits(k)]
    cmd = 1 + 2 * 9
    ref = h.oracle_render(u, cmd, words, W, H)["rgba"]
    src = rm.spec_source(cmd, words, rm.MODE_FAST)
    assert ("op_smooth2<true>(" in src) == (k <= 4.0) and ("op_smooth2<false>(" in src) == (k > 4.0)   # fast mode: rays in pairs
    ctx.set_mode(rm.MODE_FAST)
    ctx.set_kernel(rm.KERNEL_SIMPLE)
    plain = ctx.render(u, cmd, words, W, H)          # map_eval1: always max(.,0)
    ctx.set_kernel(rm.KERNEL_DIRECT)
    imgs = {}
    ctx.set_specialize(False)
    imgs["interpreter"] = ctx.render(u, cmd, words, W, H)
    for name, after in (("structure", 0), ("literal", 1)):
        ctx.set_spec_literal_after(after)
        ctx.prepare(u, cmd, words)
        ctx.specialize()
        ctx.paint(W, H)
        ctx.paint(W, H)
        assert ctx.spec_tier == (1 if name == "structure" else 2)
        imgs[name] = ctx.download(np.zeros((H, W, 4), np.uint8))
    imgs["one-sample kernel"] = plain
    for name, img in imgs.items():
        g = gates(img, ref)                          # a blend clamped wrongly would move the whole surface (silhouette
        assert g["rgba_frac"] <= 1e-2, (name, k, g)  # flips alone reach 0.6 % of the pixels of so small a frame at k = 0.05)
    for name in ("structure", "literal"):            # the production kernel's three builds: contraction differences only
        d = np.abs(imgs[name].astype(np.int32) - imgs["interpreter"].astype(np.int32)).max(-1)
        assert int(d.max()) <= 2 and float((d > 0).mean()) <= 5e-3, (name, k, int(d.max()), float((d > 0).mean()))
    ctx.set_mode(rm.MODE_EXACT)                      # and exact mode is what it was
    ctx.set_specialize(False)
    assert np.array_equal(ctx.render(u, cmd, words, W, H), ref)
