"""The device-side interpreters and the run-time generated map_scene code, compiled for the HOST and checked against
the oracle — no GPU needed.

tests/hostcheck/ compiles ray-marching_b200/csrc/rm_program.cpp (host decoder) together with the device headers
rm_device.cuh (map_eval1: verification kernels) and rm_wavefront_common.cuh (map_evalR: production kernel) through a
stand-in cuda_runtime.h whose *_rn intrinsics are plain IEEE operations.  Exact mode therefore computes the bits the GPU
computes, and what is checked here is everything between the command words and one map_scene value: validation,
lowering of the position stack (N4), operand order, spill slots, chain re-use.  It is test infrastructure: the product
library is never built this way and has no CPU path; kernel launch plumbing is covered by the `-m gpu` tests.
"""
import ctypes as C
import re
import subprocess
from pathlib import Path

import numpy as np
import pytest

import helpers as h

HC_DIR = Path(__file__).resolve().parent / "hostcheck"
CSRC = h.ROOT / "ray-marching_b200" / "csrc"
u32p, f32p = C.POINTER(C.c_uint32), C.POINTER(C.c_float)
CXX = ["g++", "-O1", "-std=c++17", "-ffp-contract=off", "-fPIC", "-shared", "-pthread", f"-I{HC_DIR / 'shim'}", f"-I{CSRC}", "-w"]

PTS = np.random.default_rng(11).uniform(-4, 4, (1003, 3)).astype(np.float32)  # not a multiple of 8: ragged last group


def load_hostcheck():
    out = HC_DIR / "_build" / "libhostcheck.so"
    out.parent.mkdir(exist_ok=True)
    deps = [HC_DIR / "hostcheck.cpp", HC_DIR / "hc_grid.h", HC_DIR / "shim" / "cuda_runtime.h"] + [CSRC / f for f in (
        "rm_program.cpp", "rm_program.h", "rm_device.cuh", "rm_wavefront_common.cuh", "rm_wavefront.h", "rm_direct_body.cuh",
        "rm_frame.h")]
    if not out.exists() or out.stat().st_mtime < max(d.stat().st_mtime for d in deps):
        subprocess.run(CXX + ["-o", str(out), str(HC_DIR / "hostcheck.cpp")], check=True, capture_output=True)
    lib = C.CDLL(str(out))
    lib.hc_map.restype = C.c_int
    lib.hc_map.argtypes = [C.c_int, C.c_int, C.c_int, C.c_uint32, u32p, C.c_uint32, C.c_float, f32p, C.c_uint64, f32p, C.c_int]
    lib.hc_decode.restype = C.c_int
    lib.hc_decode.argtypes = [C.c_uint32, u32p, C.c_uint32, C.c_uint32, C.c_void_p, C.c_uint32, u32p, C.POINTER(C.c_int), u32p, C.c_int]
    return lib


@pytest.fixture(scope="module")
def hc():
    return load_hostcheck()


def hc_map(lib, cmd, words, pts, exact=True, R=1, force_ext=False, collapse=False):
    w = np.ascontiguousarray(np.asarray(words, np.uint32))
    pts = np.ascontiguousarray(pts, np.float32)
    out = np.zeros(len(pts), np.float32)
    rc = lib.hc_map(int(exact), R, int(force_ext), cmd, w.ctypes.data_as(u32p), w.size, 100.0, pts.ctypes.data_as(f32p),
                    len(pts), out.ctypes.data_as(f32p), int(collapse))
    assert rc == 0, rc
    return out


def hc_decode(lib, cmd, words, max_words=3072, collapse=False):
    w = np.ascontiguousarray(np.asarray(words, np.uint32))
    w = w if w.size else np.zeros(1, np.uint32)
    buf = np.zeros(4096 * 12, np.uint32)
    n, ext, fl = C.c_uint32(), C.c_int(), C.c_uint32()
    rc = lib.hc_decode(cmd, w.ctypes.data_as(u32p), int(np.asarray(words).size), max_words, buf.ctypes.data_as(C.c_void_p), 4096,
                       C.byref(n), C.byref(ext), C.byref(fl), int(collapse))
    return rc, buf[:n.value * 12].reshape(-1, 12), bool(ext.value), fl.value


def same_bits(a, b):
    return np.array_equal(a.view(np.uint32), b.view(np.uint32))


# ---- the harness itself: on reference-set programs it has to reproduce what the GPU tests establish -----------------
@pytest.mark.parametrize("scene", ["S0", "S1", "S2"])
@pytest.mark.parametrize("R", [1, 4, 8])
def test_reference_set_scenes_match_the_oracle_bit_for_bit(hc, scene, R):
    cmd, words = h.SCENES[scene]()
    ref = h.oracle_map(cmd, words, PTS)
    assert same_bits(hc_map(hc, cmd, words, PTS, True, R), ref)
    assert same_bits(hc_map(hc, cmd, words, PTS, True, R, force_ext=True), ref)  # EXT interpreter, ordinary program
    np.testing.assert_allclose(hc_map(hc, cmd, words, PTS, False, R), ref, rtol=0, atol=2e-5)


@pytest.mark.parametrize("seed", range(4))
def test_reference_set_random_trees(hc, seed):
    rng = np.random.default_rng(50 + seed)
    for n_leaves in (1, 2, 3, 7, 16, 30):
        words, cmd = h.random_program(rng, n_leaves)
        ref = h.oracle_map(cmd, words, PTS)
        for R in (1, 4, 8):
            assert same_bits(hc_map(hc, cmd, words, PTS, True, R), ref)


# ---- N4: plane + space transforms ----------------------------------------------------------------------------------
@pytest.mark.parametrize("seed", range(12))
def test_extended_random_trees_match_the_oracle_bit_for_bit(hc, seed):
    rng = np.random.default_rng(700 + seed)
    n_ext = 0
    for n_leaves in (1, 2, 3, 5, 9, 14, 24):
        words, cmd = h.random_program(rng, n_leaves, ext=True)
        rc, insns, ext, _ = hc_decode(hc, cmd, words)
        assert rc == 0
        assert ext == any(op in (2, 200, 202, 204) for op in _opcodes(words))
        n_ext += ext
        ref = h.oracle_map(cmd, words, PTS)
        for R in (1, 4, 8):
            got = hc_map(hc, cmd, words, PTS, True, R)
            assert same_bits(got, ref), (seed, n_leaves, R, float(np.abs(got - ref).max()))
            fast = hc_map(hc, cmd, words, PTS, False, R)
            np.testing.assert_allclose(fast, ref, rtol=0, atol=1e-4)
    assert n_ext >= 3


def _opcodes(words):
    np_ = {0: 4, 1: 6, 2: 4, 103: 1, 200: 3, 202: 9, 204: 1, 205: 1}
    i, out = 0, []
    while i < len(words):
        out.append(int(words[i]))
        i += 1 + np_.get(int(words[i]), 0)
    return out


def test_known_cases_through_the_interpreters(hc):
    cases = [
        h.translated((1.0, -0.5, 0.25), (h.sphere((0, 0, 0), 1.0), 1)),
        h.rotated([[0, -1, 0], [1, 0, 0], [0, 0, 1]], (h.box((1, 0, 0), (0.2, 0.3, 0.4)), 1)),
        h.scaled(3.0, (h.sphere((0, 0, 0), 1.0) + h.box((0.75, 0, 0), (0.5, 0.5, 0.5)) + [h.OP_SMOOTH_UNION, h.fbits(0.25)], 3)),
        (h.plane((0.6, 0.0, 0.8), -0.25), 1),
        # a transformed right operand: evaluated in its own frame, combined with a value from the outer frame
        (h.box((0, 0, 0), (1, 1, 1)) + h.translated((3, 0, 0), (h.sphere((0, 0, 0), 1.0), 1))[0] + [h.OP_UNION], 5),
        # the chain in force shrinks and grows again: Translate(Union(A, Rotate(B)), then C outside, then Scale(D)
        (h.translated((1, 0, 0), (h.sphere((0, 0, 0), 0.5) + h.rotated(h.rotation_matrix((1, 2, 3), 0.7), (h.box((0.5, 0, 0), (0.3, 0.2, 0.1)), 1))[0]
                                  + [h.OP_UNION], 5))[0]
         + h.box((-1, 0, 0), (0.2, 0.2, 0.2)) + [h.OP_SMOOTH_UNION, h.fbits(0.3)]
         + h.scaled(1.3, (h.sphere((0, 1, 0), 0.4), 1))[0] + [h.OP_SUBTRACTION], 13),
        # 8 nested groups
        _nest8(),
    ]
    for words, cmd in cases:
        assert h.oracle_validate(cmd, words) == 0
        ref = h.oracle_map(cmd, words, PTS)
        for R in (1, 4, 8):
            assert same_bits(hc_map(hc, cmd, words, PTS, True, R), ref), (cmd, R)


def _nest8():
    node = (h.box((0.1, 0.2, 0.3), (0.5, 0.4, 0.3)), 1)
    rng = np.random.default_rng(1)
    for d in range(8):
        node = [h.translated(rng.uniform(-1, 1, 3), node), h.rotated(h.rotation_matrix(rng.normal(size=3), rng.uniform(-3, 3)), node),
                h.scaled(float(rng.uniform(0.5, 2.0)), node)][d % 3]
    return node


def test_lowering_reuses_a_chain_and_extends_a_prefix(hc):
    A, B, Cc = h.sphere((0, 0, 0), 1.0), h.box((1, 0, 0), (0.5, 0.5, 0.5)), h.sphere((0, 2, 0), 0.5)
    # Translate(Union(A, B)): one XF instruction, then two primitives that read the same q
    w, c = h.translated((1, 2, 3), (A + B + [h.OP_UNION], 3))
    rc, insns, ext, _ = hc_decode(hc, c, w)
    assert rc == 0 and ext and len(insns) == 3
    assert insns[0][9] == 32 and insns[0][10] == 1          # XOP_TRANSLATE from the ray position
    # Translate(Union(A, Scale(B))) then C: T | A | S (continues from q) | B push | ACCMUL | pop-union | RESET | C fused
    inner = (A + h.scaled(2.0, (B, 1))[0] + [h.OP_UNION], 5)
    w, c = h.translated((1, 2, 3), inner)
    w, c = w + Cc + [h.OP_UNION], c + 2
    rc, insns, ext, _ = hc_decode(hc, c, w)
    assert rc == 0
    ops = [int(i[9]) for i in insns]
    assert ops == [32, 4, 34, 11, 36, 0, 35, 6], ops
    assert [int(i[10]) & 1 for i in insns if int(i[9]) in (32, 34)] == [1, 0]


def test_documented_instruction_counts(hc):
    # DESIGN.md 6.4 / profiles/r1_s17_NOTES.md quote these
    for scene, n_insn, ext in (("S0", 1, False), ("S1", 3, False), ("S2", 32, False), ("S3", 42, True)):
        cmd, words = h.SCENES[scene]()
        rc, insns, e, _ = hc_decode(hc, cmd, words)
        assert rc == 0 and len(insns) == n_insn and e == ext, (scene, len(insns), e)


def test_instruction_stream_can_outgrow_the_word_stream(hc):
    # the reason rm_prepare checks the decoded length against the context's capacity (tests/test_n4_gpu.py)
    words, cmd = h.program_longer_than_its_words()
    rc, insns, ext, _ = hc_decode(hc, cmd, words)
    assert rc == 0 and ext and len(insns) > len(words), (len(insns), len(words))
    pts = PTS[:64]
    assert same_bits(hc_map(hc, cmd, words, pts, True, 8), h.oracle_map(cmd, words, pts))


# ---- fast-form lowering of extended programs (what RM_MODE_FAST paints): one similarity per primitive ------------------------------
@pytest.mark.parametrize("seed", range(8))
def test_collapsed_chains_agree_with_the_oracle_within_fast_mode_tolerance(hc, seed):
    rng = np.random.default_rng(1300 + seed)
    for n_leaves in (1, 2, 3, 5, 9, 14, 24):
        words, cmd = h.random_program(rng, n_leaves, ext=True)
        ref = h.oracle_map(cmd, words, PTS)
        rc, insns, ext, _ = hc_decode(hc, cmd, words, collapse=True)
        rc2, insns2, _, _ = hc_decode(hc, cmd, words)
        assert rc == 0 and rc2 == 0 and len(insns) <= len(insns2)
        for R in (1, 4, 8):
            got = hc_map(hc, cmd, words, PTS, False, R, collapse=True)
            np.testing.assert_allclose(got, ref, rtol=0, atol=1e-4)
    # the 8-deep nest collapses to ONE transform instruction + the primitive + the scale pops' multiplications
    words, cmd = _nest8()
    rc, insns, _, _ = hc_decode(hc, cmd, words, collapse=True)
    assert [int(i[9]) for i in insns][:2] == [33, 10] and all(int(i[9]) == 36 for i in insns[2:])
    np.testing.assert_allclose(hc_map(hc, cmd, words, PTS, False, 8, collapse=True), h.oracle_map(cmd, words, PTS), rtol=0, atol=1e-4)


def test_collapsed_s3_is_shorter_and_translation_only_chains_cost_nothing(hc):
    cmd, words = h.scene_s3()
    assert len(hc_decode(hc, cmd, words, collapse=True)[1]) == 29 and len(hc_decode(hc, cmd, words)[1]) == 42
    np.testing.assert_allclose(hc_map(hc, cmd, words, PTS, False, 8, collapse=True), h.oracle_map(cmd, words, PTS), rtol=0, atol=1e-4)
    # Translate(Union(A, Translate(B))): no transform instruction at all, the offsets moved into the centres
    inner = (h.sphere((0, 0, 0), 1.0) + h.translated((0, 2, 0), (h.box((1, 0, 0), (0.5, 0.5, 0.5)), 1))[0] + [h.OP_UNION], 5)
    w, c = h.translated((1, 2, 3), inner)
    rc, insns, ext, _ = hc_decode(hc, c, w, collapse=True)
    assert rc == 0 and ext and [int(i[9]) for i in insns] == [4, 12]
    assert insns[0][:3].view(np.float32).tolist() == [1, 2, 3] and insns[1][:3].view(np.float32).tolist() == [2, 4, 3]
    # without the flag nothing changes: the default lowering is the sequential one
    assert [int(i[9]) for i in hc_decode(hc, c, w)[1]] == [32, 4, 32, 12]


def test_flops_per_eval_counts_the_sequential_semantics(hc):
    w, c = h.translated((1, 2, 3), (h.sphere((0, 0, 0), 1) + h.plane((0, 1, 0), 1) + [h.OP_UNION], 3))
    w, c = h.scaled(2.0, h.rotated(np.eye(3), (w, c)))
    _, _, _, flops = hc_decode(hc, c, w)
    assert flops == 10 + 6 + 1 + 3 + 15 + 4


@pytest.mark.parametrize("words,count,code", [
    ([h.OP_TRANSLATION_POP], 1, -5),
    ([h.OP_TRANSLATION_PUSH, 0, 0, 0] + h.sphere((0, 0, 0), 1), 2, -5),
    ([h.OP_TRANSLATION_PUSH, 0, 0, 0] + h.sphere((0, 0, 0), 1) + [h.OP_ROTATION_POP], 3, -5),
    ([h.OP_TRANSLATION_PUSH, 0, 0, 0, h.OP_TRANSLATION_POP], 2, -5),
    (h.sphere((0, 0, 0), 1) + [h.OP_TRANSLATION_PUSH, 0, 0, 0] + h.sphere((0, 0, 0), 1) + [h.OP_UNION, h.OP_TRANSLATION_POP], 5, -5),
    ([h.OP_SCALE_PUSH, h.fbits(2.0)] + h.sphere((0, 0, 0), 1) + [h.OP_SCALE_POP, h.fbits(3.0)], 3, -8),
    ([h.OP_SCALE_PUSH, h.fbits(0.0)] + h.sphere((0, 0, 0), 1) + [h.OP_SCALE_POP, h.fbits(0.0)], 3, -8),
    ([h.OP_ROTATION_PUSH] + [0] * 8, 1, -5),
    ([h.OP_TRANSLATION_PUSH, 0, 0, 0] * 9 + h.sphere((0, 0, 0), 1) + [h.OP_TRANSLATION_POP] * 9, 19, -4),
    ([h.OP_TRANSLATION_PUSH, 0, 0, 0] * 8 + h.sphere((0, 0, 0), 1) + [h.OP_TRANSLATION_POP] * 8, 17, 0),
])
def test_validation_codes_for_transform_groups(hc, words, count, code):
    assert hc_decode(hc, count, words)[0] == code


def test_validation_agrees_with_the_oracle_on_random_extended_streams(hc):
    rng = np.random.default_rng(2)
    ops = [0, 1, 2, 100, 101, 103, 200, 201, 202, 203, 204, 205, 7]
    npar = {0: 4, 1: 6, 2: 4, 103: 1, 200: 3, 202: 9, 204: 1, 205: 1}
    agree = 0
    for _ in range(6000):
        n = int(rng.integers(1, 10))
        words = []
        for _ in range(n):
            t = ops[rng.integers(len(ops))]
            words.append(t)
            words += [h.fbits([0.5, 2.0][rng.integers(2)])] * npar.get(t, 0)
        a = hc_decode(hc, n, words)[0] == 0
        b = h.oracle_validate(n, words) == 0
        assert a == b, words
        agree += a
    # valid ones also come from the generator, mutated
    for _ in range(600):
        words, cmd = h.random_program(rng, int(rng.integers(1, 8)), ext=True)
        if rng.random() < 0.5:  # replace one opcode word (parameters stay valid numbers)
            pos, i = [], 0
            while i < len(words):
                pos.append(i)
                i += 1 + npar.get(int(words[i]), 0)
            words[pos[rng.integers(len(pos))]] = ops[rng.integers(len(ops))]
        a = hc_decode(hc, cmd, words)[0] == 0
        b = h.oracle_validate(cmd, words) == 0
        assert a == b, words
        agree += a
    assert agree > 300


# ---- run-time generated map_scene code (rm_jit.cu) ------------------------------------------------------------------
def _spec_params(rm_lib, cmd, w, mode_exact):
    """The parameter block the generated structure-tier kernel reads (rm_spec_params), padded like the kernel argument."""
    rm_lib.rm_spec_params.restype = C.c_int
    rm_lib.rm_spec_params.argtypes = [C.c_uint32, u32p, C.c_uint32, C.c_int, f32p, C.c_size_t, C.POINTER(C.c_size_t)]
    need = C.c_size_t()
    rc = rm_lib.rm_spec_params(cmd, w.ctypes.data_as(u32p), w.size, 0 if mode_exact else 1, None, 0, C.byref(need))
    assert rc == 0, rc
    out = np.zeros(max(4, need.value), np.float32)
    rm_lib.rm_spec_params(cmd, w.ctypes.data_as(u32p), w.size, 0 if mode_exact else 1, out.ctypes.data_as(f32p), out.size, None)
    return out


def _spec_map_on_host(tmp_path, rm_lib, cmd, words, mode_exact, R, pts):
    """Compiles the generated SpecMap functor for the host (kernel wrapper cut off) and evaluates it at pts."""
    w = np.ascontiguousarray(np.asarray(words, np.uint32))
    buf = C.create_string_buffer(1 << 20)
    need = C.c_size_t()
    rc = rm_lib.rm_spec_source(cmd, w.ctypes.data_as(u32p), w.size, 0 if mode_exact else 1, R, buf, 1 << 20, C.byref(need))
    assert rc == 0, rc
    src = buf.value.decode()
    m = re.search(r"rm::SpecMap<(true|false), (\d)> map\(", src)
    ex, r = m.group(1), int(m.group(2))
    src = src[:src.index('extern "C" __global__')]
    params = _spec_params(rm_lib, cmd, w, mode_exact)
    src = '#include "cuda_runtime.h"\n' + src + f"""
extern "C" void hc_spec(const float *pts, unsigned long long n, float max_dist, const float *params, float *out) {{
    rm::SpecParams P;
    memcpy(&P, params, sizeof P);
    rm::SpecMap<{ex}, {r}> map(max_dist, P);
    for (unsigned long long i = 0; i < n; i += {r}) {{
        float px[{r}], py[{r}], pz[{r}], s[{r}];
        for (int k = 0; k < {r}; ++k) {{ unsigned long long j = i + k < n ? i + k : n - 1; px[k] = pts[3 * j]; py[k] = pts[3 * j + 1]; pz[k] = pts[3 * j + 2]; }}
        map(px, py, pz, s);
        for (int k = 0; k < {r} && i + k < n; ++k) out[i + k] = s[k];
    }}
}}
"""
    # a fresh file name per build: dlopen of a path it has already loaded would hand back the OLD code
    tag = f"{abs(hash((cmd, bytes(w), mode_exact, R)))}"
    cu = tmp_path / f"spec_{tag}.cpp"
    cu.write_text(src)
    so = tmp_path / f"libspec_{tag}.so"
    subprocess.run(CXX + ["-o", str(so), str(cu)], check=True, capture_output=True)
    lib = C.CDLL(str(so))
    lib.hc_spec.argtypes = [f32p, C.c_uint64, C.c_float, f32p, f32p]
    pts = np.ascontiguousarray(pts, np.float32)
    out = np.zeros(len(pts), np.float32)
    lib.hc_spec(pts.ctypes.data_as(f32p), len(pts), 100.0, params.ctypes.data_as(f32p), out.ctypes.data_as(f32p))
    return out, r


@pytest.fixture(scope="module")
def rm_lib():
    so = h.ROOT / "ray-marching_b200" / "librm_b200.so"
    if not so.exists():
        pytest.skip("librm_b200.so not built")
    lib = C.CDLL(str(so))
    lib.rm_spec_source.restype = C.c_int
    lib.rm_spec_source.argtypes = [C.c_uint32, u32p, C.c_uint32, C.c_int, C.c_int, C.c_char_p, C.c_size_t, C.POINTER(C.c_size_t)]
    return lib


@pytest.mark.parametrize("seed", range(5))
def test_generated_code_of_extended_programs_matches_the_oracle(hc, rm_lib, tmp_path, seed):
    rng = np.random.default_rng(900 + seed)
    words, cmd = h.random_program(rng, [3, 6, 10, 15, 2][seed], ext=True)
    if seed == 0:  # make sure every N4 command is in at least one compiled program
        words, cmd = h.scaled(1.5, h.rotated(h.rotation_matrix((0, 1, 1), 0.9), h.translated((0.5, 0, 0), (words + h.plane((0, 1, 0), 1.0)
                                                                                                               + [h.OP_UNION], cmd + 2))))
    ref = h.oracle_map(cmd, words, PTS)
    got, r = _spec_map_on_host(tmp_path, rm_lib, cmd, words, True, 0, PTS)
    assert same_bits(got, ref), (seed, r, float(np.abs(got - ref).max()))
    fast, _ = _spec_map_on_host(tmp_path, rm_lib, cmd, words, False, 0, PTS)
    np.testing.assert_allclose(fast, ref, rtol=0, atol=1e-4)


def test_generated_code_of_the_reference_set_matches_too(hc, rm_lib, tmp_path):
    cmd, words = h.scene_s1()
    got, _ = _spec_map_on_host(tmp_path, rm_lib, cmd, words, True, 0, PTS)
    assert same_bits(got, h.oracle_map(cmd, words, PTS))


def test_epilogue_constants_are_the_correctly_rounded_values_the_shader_computes():
    """rm_device.cuh holds the four gamma'd floor colours as constants and divides the 16-sample sum by 16 with a multiply:
    both must be what the oracle's operation-by-operation arithmetic gives (wgsl:124-127, 68, 75)."""
    import re
    src = (CSRC / "rm_device.cuh").read_text()
    got = {m.group(1): np.float32(m.group(2)) for m in re.finditer(r"#define RM_GAMMA_FLOOR_(\w+) ([0-9.]+)f", src)}
    f = np.float32
    want = {"DARK_RG": np.sqrt(f(0.1) + f(0.2) * f(0)), "DARK_B": np.sqrt(f(0.2) + f(0.2) * f(0)),
            "LIGHT_RG": np.sqrt(f(0.1) + f(0.2) * f(1)), "LIGHT_B": np.sqrt(f(0.2) + f(0.2) * f(1))}
    assert set(got) == set(want)
    for k in want:
        assert want[k].dtype == np.float32 and got[k].view(np.uint32) == want[k].view(np.uint32), k
    assert f(0.1) + f(0.2) * f(1) > f(0.2) and not f(0.1) + f(0.2) * f(0) > f(0.2)   # the light / dark test of the code
    rng = np.random.default_rng(5)
    x = np.concatenate([rng.integers(0, 2**32, 2_000_000, dtype=np.uint64).astype(np.uint32).view(np.float32),
                        np.array([0.0, -0.0, 1e-45, 2e-44, 1.7e-38, 16.0, 3.4e38, np.inf, -np.inf], np.float32)])
    with np.errstate(all="ignore"):
        a, b = x / f(16.0), x * f(0.0625)
    ok = (a.view(np.uint32) == b.view(np.uint32)) | (np.isnan(a) & np.isnan(b))
    assert ok.all()
