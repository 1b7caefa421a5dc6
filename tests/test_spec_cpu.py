"""Program specialisation, the parts that need no GPU: the source generator (straight-line map_scene from the postfix
command stream) and a trial NVRTC compile for sm_100a (NVRTC cross-compiles without a device, like nvcc)."""
import re

import numpy as np
import pytest
from pathlib import Path

import helpers as h
import rm_import


@pytest.fixture(scope="module")
def rm():
    m = rm_import.load()
    if not m.LIB_PATH.exists():
        m.build()
    return m


def _defs(src):
    return re.findall(r"RM_FOR_R (v\d+)\[r\] = (\w+)", src)


def test_source_is_one_ssa_value_per_command_in_stream_order(rm, monkeypatch):
    monkeypatch.setenv("RM_SPEC_PACKED", "0")   # the scalar form of the fast-mode source (the packed one: test_fast_mode_source_evaluates_rays_in_pairs)
    cmd, words = h.scene_s1()  # box, sphere, union, sphere, subtraction  (main.rs:112-126)
    src = rm.spec_source(cmd, words, rm.MODE_FAST)
    assert [f for _, f in _defs(src)] == ["sd_box", "sd_sphere", "fminf", "sd_sphere", "fmaxf"]
    assert "v2[r] = fminf(v0[r], v1[r])" in src              # a = deeper entry, b = top (wgsl:243-244)
    assert "v4[r] = fmaxf(v2[r], -v3[r])" in src             # A \ B = max(a, -b)     (wgsl:248-252)
    assert "RM_FOR_R out[r] = v4[r];" in src
    assert 'extern "C" __global__' in src and "rm::direct_body<false, 4, false, false>" in src
    assert "direct_body<true, 4, false, false>" in rm.spec_source(cmd, words, rm.MODE_EXACT)


def test_literal_tier_emits_constants_by_their_bits(rm, monkeypatch):
    monkeypatch.setenv("RM_SPEC_PACKED", "0")   # the scalar form of the fast-mode source (the packed one: test_fast_mode_source_evaluates_rays_in_pairs)
    lit = rm.MODE_FAST | rm.SPEC_LITERAL_TIER
    words = h.sphere((0.1, -2.5, 3.0e-7), 1.25)
    src = rm.spec_source(1, words, lit)
    for w in words[1:]:
        assert "__uint_as_float(0x%08xu)" % w in src
    k = 0.3
    words = h.sphere((0, 0, 0), 1) + h.sphere((1, 0, 0), 1) + [103, h.fbits(k)]
    # exact mode reads k (and 1/k when it is exact); fast mode reads k*s and s with s = sqrt(1/(4k))
    src = rm.spec_source(3, words, rm.MODE_EXACT | rm.SPEC_LITERAL_TIER)
    inv = np.float32(1.0) / np.float32(k)
    c4k = np.float32(0.25) / np.float32(k)
    m = re.search(r"op_smooth<EX>\(v0\[r\], v1\[r\], (\S+), (\S+), (\S+), (true|false)\)", src)
    assert m and m.group(1) == "__uint_as_float(0x%08xu)" % h.fbits(k)
    assert m.group(2) == "__uint_as_float(0x%08xu)" % int(inv.view(np.uint32))
    assert m.group(3) == "__uint_as_float(0x%08xu)" % int(c4k.view(np.uint32))
    assert m.group(4) == "false"
    src = rm.spec_source(3, words, lit)
    s = np.float32(np.sqrt(0.25 / np.float64(np.float32(k))))
    ks = np.float32(np.float64(np.float32(k)) * np.sqrt(0.25 / np.float64(np.float32(k))))
    # k*s <= 1 (k <= 4): the clamp at 0 rides on the FFMA as its .SAT modifier (op_smooth<EX, true>)
    m = re.search(r"op_smooth<EX, true>\(v0\[r\], v1\[r\], 0.0f, (\S+), (\S+), false\)", src)
    assert m and m.group(1) == "__uint_as_float(0x%08xu)" % int(ks.view(np.uint32)) and m.group(2) == "__uint_as_float(0x%08xu)" % int(s.view(np.uint32))
    big = list(words)
    big[-1] = h.fbits(4.5)      # k*s = sqrt(k)/2 > 1: the general form
    assert "op_smooth<EX, false>(" in rm.spec_source(3, big, lit) and "op_smooth<EX, true>(" not in rm.spec_source(3, big, lit)
    four = list(words)
    four[-1] = h.fbits(4.0)     # k*s = 1 exactly: still a saturate
    assert "op_smooth<EX, true>(" in rm.spec_source(3, four, lit)
    # structure tier: which of the two forms is structural, the value of k is not
    st = lambda kk: rm.spec_source(3, words[:-1] + [h.fbits(kk)], rm.MODE_FAST)
    assert st(0.3) == st(3.9) and st(0.3) != st(4.5) and st(4.5) == st(100.0)
    words[-1] = h.fbits(0.25)  # power of two: the exact-mode divide becomes a multiply by the exact reciprocal
    assert ", true)" in rm.spec_source(3, words, rm.MODE_EXACT | rm.SPEC_LITERAL_TIER)


def test_structure_tier_reads_parameters_from_the_kernel_argument(rm, monkeypatch):
    monkeypatch.setenv("RM_SPEC_PACKED", "0")   # the scalar form of the fast-mode source (the packed one: test_fast_mode_source_evaluates_rays_in_pairs)
    """The default kernel depends on the STRUCTURE only: parameters are P.f[i] (a __grid_constant__ kernel argument), except
    exact zeros, which stay literals (and are part of the structure)."""
    a = h.sphere((0.1, -2.5, 0.0), 1.25) + h.box((0, 0, 0), (1, 2, 3)) + [103, h.fbits(0.3)]
    b = h.sphere((0.7, 9.5, 0.0), 0.25) + h.box((0, 0, 0), (3, 1, 2)) + [103, h.fbits(0.7)]
    sa, sb = rm.spec_source(3, a), rm.spec_source(3, b)
    assert sa == sb                                     # same structure, other values: the same code
    assert "__uint_as_float" not in sa and "const __grid_constant__ rm::SpecParams P" in sa
    assert re.search(r"sd_sphere<EX>\(px\[r\], py\[r\], pz\[r\], P.f\[0\], P.f\[1\], 0.0f, P.f\[2\]\)", sa)
    assert re.search(r"sd_box<EX>\(px\[r\], py\[r\], pz\[r\], 0.0f, 0.0f, 0.0f, P.f\[4\], P.f\[5\], P.f\[6\]\)", sa)   # 16-byte aligned group
    pa, pb = rm.spec_params(3, a), rm.spec_params(3, b)
    assert pa.size == pb.size and pa.size % 4 == 0
    np.testing.assert_array_equal(pa[:3], np.float32([0.1, -2.5, 1.25]))
    np.testing.assert_array_equal(pa[4:7], np.float32([1, 2, 3]))
    np.testing.assert_array_equal(pb[:3], np.float32([0.7, 9.5, 0.25]))
    c = h.sphere((0.1, -2.5, 0.5), 1.25) + h.box((0, 0, 0), (1, 2, 3)) + [103, h.fbits(0.3)]
    assert rm.spec_source(3, c) != sa                   # a zero became non-zero: another structure
    # exact mode: whether k is a power of two is structural (multiply by 1/k instead of dividing)
    e1 = rm.spec_source(3, a, rm.MODE_EXACT)
    a2 = list(a)
    a2[-1] = h.fbits(0.25)
    assert rm.spec_source(3, a2, rm.MODE_EXACT) != e1 and rm.spec_source(3, a2, rm.MODE_FAST) == sa


def test_rays_per_thread_follow_the_code_size(rm):
    cmd, words = h.scene_s0()
    assert "direct_body<false, 4," in rm.spec_source(cmd, words)
    cmd, words = h.scene_s2()  # 63 commands: ~9 KB of code per ray -> 2 rays keep the march loop in the instruction cache
    src = rm.spec_source(cmd, words)
    # structure tier of a deep program: ONE out-of-line copy of the scene code too, its parameters in the module's
    # __constant__ block (an out-of-line function cannot read the kernel's arguments as constants)
    assert "direct_body<false, 2," in src and "__launch_bounds__(RM_DIRECT_THREADS, 10)" in src and "__noinline__ Out eval(" in src
    assert "__constant__ rm::SpecParams g_rm_spec_P;" in src and "g_rm_spec_P.f[0]" in src and " P.f[" not in src and "@P@" not in src
    s1 = rm.spec_source(*h.scene_s1())                    # shallow: inlined, parameters from the kernel argument
    assert "__constant__" not in s1 and "P.f[0]" in s1 and "__noinline__" not in s1 and "@P@" not in s1
    lit = rm.spec_source(cmd, words, rm.MODE_FAST | rm.SPEC_LITERAL_TIER)   # the literal tier of a deep program: ONE out-of-line copy of the scene code
    assert "direct_body<false, 2," in lit and "__launch_bounds__(RM_DIRECT_THREADS, 11)" in lit   # in ray pairs: 11 resident blocks (10 scalar)
    assert "__launch_bounds__(RM_DIRECT_THREADS, 10)" in rm.spec_source(cmd, words, rm.MODE_EXACT | rm.SPEC_LITERAL_TIER)
    assert "__noinline__ Out eval(float x0, float y0, float z0, float x1, float y1, float z1)" in lit
    assert "direct_body<false, 8," in rm.spec_source(cmd, words, rm.MODE_FAST, 8)  # explicit request is honoured
    # a chain far beyond the instruction-cache budget is refused (the interpreter is the better kernel there)
    n = 120
    long_words = h.box((0, 0, 0), (1, 1, 1))
    for _ in range(n - 1):
        long_words += h.box((0.5, 0, 0), (1, 1, 1)) + [103, h.fbits(0.25)]
    with pytest.raises(rm.RmError) as e:
        rm.spec_source(2 * n - 1, long_words)
    assert e.value.code == -9


def test_deep_stacks_and_empty_program(rm, monkeypatch):
    monkeypatch.setenv("RM_SPEC_PACKED", "0")   # the scalar form of the fast-mode source (the packed one: test_fast_mode_source_evaluates_rays_in_pairs)
    # ((s0 u s1) n (s2 \ s3)): right operand is a subtree -> the interpreter needs a spill slot, SSA does not care
    s = [h.sphere((i, 0, 0), 1) for i in range(4)]
    words = s[0] + s[1] + [100] + s[2] + s[3] + [101] + [102]
    src = rm.spec_source(7, words)
    assert "v6[r] = fmaxf(v2[r], v5[r])" in src and "v5[r] = fmaxf(v3[r], -v4[r])" in src
    src = rm.spec_source(0, [])
    assert "out[r] = max_dist" in src


@pytest.mark.parametrize("bad,code", [([5], -3), ([100], -5), (h.sphere((0, 0, 0), 1)[:3], -5)])
def test_invalid_streams_are_rejected_with_the_validation_codes(rm, bad, code):
    with pytest.raises(rm.RmError) as e:
        rm.spec_source(1, bad)
    assert e.value.code == code


def test_generator_accepts_what_the_decoder_accepts(rm, monkeypatch):
    monkeypatch.setenv("RM_SPEC_PACKED", "0")   # the scalar form of the fast-mode source (the packed one: test_fast_mode_source_evaluates_rays_in_pairs)
    """ADVICE r1: a stream that leaves MORE than one value is valid — the shader's final pop_value takes the top
    (wgsl:200-202) — for rm_prepare and for the specialiser alike."""
    two = h.sphere((0, 0, 0), 1) + h.sphere((2, 0, 0), 0.5)
    assert rm.validate_program(2, two) == 0
    src = rm.spec_source(2, two)
    assert "RM_FOR_R out[r] = v1[r];" in src


@pytest.mark.parametrize("scene", ["s0", "s1", "s2"])
@pytest.mark.parametrize("mode", [0, 1, 0x100, 0x101])
def test_generated_source_compiles_for_sm_100a(rm, scene, mode):
    cmd, words = getattr(h, "scene_" + scene)()
    r = rm.spec_compile_check(cmd, words, mode)
    if r["rc"] == -9 and "NVRTC is not available" in r["log"]:
        pytest.skip("no libnvrtc in this environment: " + r["log"][:200])
    assert r["rc"] == 0, r["log"]
    assert r["cubin_bytes"] > 10000 and r["compile_ms"] > 0


def test_extended_program_generates_position_ssa_and_compiles_for_sm_100a(rm):
    """N4: the evaluation point is an SSA value too (q<i>x/y/z defined by the Push, dropped by the Pop)."""
    inner = (h.sphere((0, 0, 0), 1.0) + h.plane((0, 1, 0), 0.5) + [103, h.fbits(0.25)], 3)
    words, cmd = h.scaled(2.0, h.rotated(h.rotation_matrix((0, 0, 1), 0.5), h.translated((1, 0, 0), inner)))
    words, cmd = words + h.box((0, 0, 0), (1, 1, 1)) + [100], cmd + 2
    src = rm.spec_source(cmd, words)
    assert "xf_scale<EX>(px[r], py[r], pz[r]," in src and "q0x[r], q0y[r], q0z[r])" in src
    assert "xf_rotate<EX>(q0x[r], q0y[r], q0z[r]," in src and "xf_translate<EX>(q1x[r], q1y[r], q1z[r]," in src
    assert "sd_sphere<EX>(q2x[r], q2y[r], q2z[r]," in src and "sd_plane<EX>(q2x[r], q2y[r], q2z[r]," in src
    assert "v8[r] = Arith<EX>::mul(v5[r]," in src          # ScalePop (command 8) rescales the group's value
    assert "sd_box<EX>(px[r], py[r], pz[r]," in src         # after the pops the root point is in force again
    for mode in (0, 1):
        r = rm.spec_compile_check(cmd, words, mode)
        if r["rc"] == -9 and "NVRTC is not available" in r["log"]:
            pytest.skip("no libnvrtc in this environment: " + r["log"][:200])
        assert r["rc"] == 0, r["log"]


def test_fast_mode_source_evaluates_rays_in_pairs(rm):
    """Fast mode, reference-set programs: the scene code works on PAIRS of rays (f2_t: FADD2 / FMUL2 / FFMA2 of sm_100 issue
    once for two rays), same operations per lane; exact mode and programs with N4 commands keep the scalar form."""
    cmd, words = h.scene_s1()
    src = rm.spec_source(cmd, words, rm.MODE_FAST)
    assert re.findall(r"RM_FOR_H (w\d+)\[h\] = (\w+)", src) == [("w0", "sd_box2"), ("w1", "sd_sphere2"), ("w2", "f2_min"), ("w3", "sd_sphere2"), ("w4", "f2_max")]
    assert "w4[h] = f2_max(w2[h], f2_neg(w3[h]))" in src and "PX[h] = f2_pack(px[2 * h], px[2 * h + 1])" in src
    assert "out[2 * h] = f2_lo(w4[h]); out[2 * h + 1] = f2_hi(w4[h]);" in src and "RM_FOR_R v" not in src
    ex = rm.spec_source(cmd, words, rm.MODE_EXACT)
    assert "f2_t" not in ex and "RM_FOR_R v4[r] = fmaxf(v2[r], -v3[r])" in ex
    c3, w3 = h.scene_s3()
    assert "f2_t" not in rm.spec_source(c3, w3, rm.MODE_FAST)           # N4 commands: scalar
    smooth = h.sphere((0, 0, 0), 1.0) + h.box((0.5, 0, 0), (1, 1, 1)) + [103, h.fbits(0.3)]
    assert "w2[h] = op_smooth2<true>(w0[h], w1[h], " in rm.spec_source(3, smooth, rm.MODE_FAST)
    big = smooth[:-1] + [h.fbits(9.0)]
    assert "op_smooth2<false>(" in rm.spec_source(3, big, rm.MODE_FAST)
    assert "RM_FOR_R out[r] = max_dist;" in rm.spec_source(0, [], rm.MODE_FAST)   # the empty program has nothing to pair


def test_the_newest_nvrtc_is_used_even_when_the_process_already_holds_an_older_one():
    """A process that imported torch holds torch's bundled libnvrtc.so.12 (12.8 here) and a bare soname resolves to it; 12.8
    builds the ray-pair scene code 24 % slower than the toolkit's 12.9.  rm_jit.cu loads every candidate
    and keeps the newest: with the old one preloaded the cubin must be the one the toolkit's compiler produces."""
    import os
    import subprocess
    import sys
    old = Path(sys.prefix) / "lib" / f"python{sys.version_info.major}.{sys.version_info.minor}" / "site-packages" / "nvidia" / "cuda_nvrtc" / "lib" / "libnvrtc.so.12"
    new = Path("/usr/local/cuda/lib64/libnvrtc.so.12")
    if not old.exists() or not new.exists():
        pytest.skip("needs both torch's bundled NVRTC and the toolkit's")
    code = ("import ctypes, sys\n"
            "pre = sys.argv[1]\n"
            "if pre: ctypes.CDLL(pre, mode=ctypes.RTLD_GLOBAL)\n"
            f"sys.path.insert(0, {str(h.ROOT / 'tests')!r}); sys.path.insert(0, {str(h.ROOT)!r})\n"
            "import rm_import, helpers as h\n"
            "rm = rm_import.load()\n"
            "cmd, words = h.scene_s2()\n"
            "r = rm.spec_compile_check(cmd, words, rm.MODE_FAST | rm.SPEC_LITERAL_TIER)\n"
            "assert r['rc'] == 0, r\n"
            "print(r['cubin_bytes'])\n")

    def cubin_bytes(preload, env_path=None):
        env = {k: v for k, v in os.environ.items() if k != "RM_NVRTC_PATH"}
        if env_path:
            env["RM_NVRTC_PATH"] = env_path
        r = subprocess.run([sys.executable, "-c", code, preload], capture_output=True, text=True, env=env, timeout=600)
        assert r.returncode == 0, r.stderr[-2000:]
        return int(r.stdout.strip().splitlines()[-1])

    with_new = cubin_bytes("", str(new))
    with_old = cubin_bytes("", str(old))
    assert cubin_bytes(str(old)) == with_new          # old one preloaded, no override: the newest candidate wins
    if with_old == with_new:
        pytest.skip("the two compilers produce the same cubin here: nothing to tell apart")
