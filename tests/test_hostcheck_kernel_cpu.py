"""The PRODUCTION KERNEL BODY on the host, lane for lane, against the oracle — no GPU needed.

tests/hostcheck compiles ray-marching_b200/csrc/rm_direct_body.cuh (ray generation, lock-step march of 4 or 8 sample
rays per thread, tetrahedron normals, in-order 16-sample resolve through the shuffle chain, 128-bit stores, strip / band
row mapping, work counters, AOVs) for the host and runs it with ONE HOST THREAD PER LANE: 32 threads form a warp and
meet at a barrier inside every __shfl_sync / __reduce_add_sync (tests/hostcheck/shim, hc_grid.h).  The map_scene functor is
the decoded-program interpreter (as in rm_direct.cu) or the run-time generated straight-line code (the `rm_spec` kernel
function that rm_jit.cu emits, compiled as it is).  Per-frame constants come from the same rm_frame.h the library uses.
Exact mode computes the GPU's bits, so the frames must equal the oracle's byte for byte.

Test infrastructure only: the product library is never built like this and has no CPU path; what remains GPU-only is the
launch plumbing of rm_abi.cu and the hardware itself, covered by the `-m gpu` tests.
"""
import ctypes as C
import re
import subprocess
from pathlib import Path

import numpy as np
import pytest

import helpers as h
import test_hostcheck_cpu as hcpu
from test_hostcheck_cpu import CSRC, CXX, HC_DIR, u32p

ROOT = Path(__file__).resolve().parent.parent
PAINT_ARGS = [C.c_void_p, C.c_void_p] + [C.c_uint32] * 6 + [C.c_int, C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p, C.c_void_p]


@pytest.fixture(scope="module")
def hc():
    lib = hcpu.load_hostcheck()
    lib.hc_paint.restype = C.c_int
    lib.hc_paint.argtypes = [C.c_int, C.c_int, C.c_uint32, u32p, C.c_uint32] + PAINT_ARGS
    return lib


def _call(fn, head, u, W, H, row0, nrows, strip_rows, strip_pitch, dst_frame_rows, frame_rows, limits, counters, aov):
    pitch = (W * 4 + 15) & ~15
    rows = frame_rows if dst_frame_rows else nrows
    dst = np.zeros((rows, pitch), np.uint8)
    cnt = np.zeros(4, np.uint64) if counters else None
    cls = np.zeros((nrows, W, 16), np.uint8) if aov else None
    t = np.zeros((nrows, W, 16), np.float32) if aov else None
    ub = C.create_string_buffer(u, 144)
    lb = C.create_string_buffer(limits or h.limits_bytes(), 12)
    rc = fn(*head, ub, lb, W, H, row0, nrows, strip_rows, strip_pitch, int(dst_frame_rows), dst.ctypes.data, pitch,
            cnt.ctypes.data if counters else None, cls.ctypes.data if aov else None, t.ctypes.data if aov else None)
    assert rc == 0, rc
    out = {"rgba": dst[:, :W * 4].reshape(rows, W, 4)}
    if counters:
        out["counters"] = cnt
    if aov:
        out["cls"], out["t"] = cls, t
    return out


def paint(lib, cmd, words, u, W, H, exact=True, R=0, row0=0, nrows=None, strip_rows=0, strip_pitch=0, dst_frame_rows=False,
          limits=None, counters=False, aov=False):
    nrows = H - row0 if nrows is None else nrows
    w = np.ascontiguousarray(np.asarray(words, np.uint32))
    w = w if w.size else np.zeros(1, np.uint32)
    head = (int(exact), R, cmd, w.ctypes.data_as(u32p), int(np.asarray(words).size))
    return _call(lib.hc_paint, head, u, W, H, row0, nrows, strip_rows, strip_pitch, dst_frame_rows, H, limits, counters, aov)


CASES = [("S0", 40, 30, 0.0, 0.0), ("S1", 48, 27, 0.6, 0.4), ("S2", 37, 21, 0.6, 0.4), ("S3", 48, 27, 0.6, 0.4)]


@pytest.mark.parametrize("name,W,H,yaw,pitch", CASES)
@pytest.mark.parametrize("R", [4, 8])
def test_kernel_body_exact_mode_is_bit_identical_to_the_oracle(hc, name, W, H, yaw, pitch, R):
    cmd, words = h.SCENES[name]()
    u = h.orbit_uniforms(W, H, yaw=yaw, pitch=pitch)
    ref = h.oracle_render(u, cmd, words, W, H, aov=True)
    got = paint(hc, cmd, words, u, W, H, R=R, counters=True)
    assert np.array_equal(got["rgba"], ref["rgba"])
    assert np.array_equal(got["counters"], ref["counters"])
    got = paint(hc, cmd, words, u, W, H, R=R, aov=True)
    assert np.array_equal(got["rgba"], ref["rgba"]) and np.array_equal(got["cls"], ref["cls"])
    assert np.array_equal(got["t"].view(np.uint32), ref["t"].view(np.uint32))


@pytest.mark.parametrize("path", sorted((ROOT / "tests" / "golden").glob("**/*.npz")), ids=lambda p: p.stem)
def test_kernel_body_reproduces_the_golden_frames(hc, path):
    z = np.load(path)
    g = {k: z[k] for k in z.files}
    W, H = int(g["W"]), int(g["H"])
    if W * H > 3000:  # the big fixtures: a band of 12 rows through the middle is enough here
        row0, nrows = H // 2 - 6, 12
    else:
        row0, nrows = 0, H
    got = paint(hc, int(g["cmd_count"]), g["words"], g["uniforms"].tobytes(), W, H, row0=row0, nrows=nrows,
                limits=g["limits"].tobytes(), aov=True)
    assert np.array_equal(got["rgba"], g["rgba"][row0:row0 + nrows])
    assert np.array_equal(got["cls"], g["cls"][row0:row0 + nrows])
    assert np.array_equal(got["t"].view(np.uint32), g["t"][row0:row0 + nrows].view(np.uint32))


@pytest.mark.parametrize("W,H", [(1, 1), (7, 3), (9, 5), (17, 2), (33, 9)])
def test_ragged_sizes(hc, W, H):
    cmd, words = h.scene_s1()
    u = h.orbit_uniforms(W, H, yaw=0.6, pitch=0.4)
    ref = h.oracle_render(u, cmd, words, W, H)
    for R in (4, 8):
        got = paint(hc, cmd, words, u, W, H, R=R, counters=True)
        assert np.array_equal(got["rgba"], ref["rgba"]) and np.array_equal(got["counters"], ref["counters"])


def test_bands_and_interleaved_strips_equal_the_rows_of_the_full_frame(hc):
    cmd, words = h.scene_s1()
    W, H = 24, 29
    u = h.orbit_uniforms(W, H, yaw=1.1, pitch=0.5)
    full = h.oracle_render(u, cmd, words, W, H)["rgba"]
    for row0, nrows in ((0, 5), (5, 13), (18, 11), (28, 1)):
        assert np.array_equal(paint(hc, cmd, words, u, W, H, row0=row0, nrows=nrows)["rgba"], full[row0:row0 + nrows])
    # interleaved strips of 4 rows over 3 shards (rm_paint_strips): shard s paints strips s, s+3, ... packed
    strip, world = 4, 3
    n_strips = (H + strip - 1) // strip
    for s in range(world):
        rows = [r for k in range(s, n_strips, world) for r in range(k * strip, min(H, (k + 1) * strip))]
        got = paint(hc, cmd, words, u, W, H, row0=s * strip, nrows=len(rows), strip_rows=strip, strip_pitch=strip * world)
        assert np.array_equal(got["rgba"], full[rows]), s
    # the same strips stored at their frame rows of ONE frame buffer (rm_paint_strips_into)
    frame = np.zeros_like(full)
    for s in range(world):
        rows = [r for k in range(s, n_strips, world) for r in range(k * strip, min(H, (k + 1) * strip))]
        got = paint(hc, cmd, words, u, W, H, row0=s * strip, nrows=len(rows), strip_rows=strip, strip_pitch=strip * world,
                    dst_frame_rows=True)["rgba"]
        frame[rows] = got[rows]
        untouched = np.ones(H, bool)
        untouched[rows] = False
        assert not got[untouched].any()
    assert np.array_equal(frame, full)


def test_limits_empty_scene_and_zero_iterations(hc):
    W, H = 20, 12
    u = h.orbit_uniforms(W, H, yaw=2.2, pitch=-0.3)
    cmd, words = h.scene_s2()
    for lim in (h.limits_bytes(0.02, 40.0, 24), h.limits_bytes(0.01, 100.0, 1), h.limits_bytes(0.01, 100.0, 0), h.limits_bytes(0.5, 3.0, 100)):
        ref = h.oracle_render(u, cmd, words, W, H, limits=lim)
        got = paint(hc, cmd, words, u, W, H, limits=lim, counters=True)
        assert np.array_equal(got["rgba"], ref["rgba"]) and np.array_equal(got["counters"], ref["counters"])
    ref = h.oracle_render(u, 0, [], W, H)
    got = paint(hc, 0, [], u, W, H, counters=True)
    assert np.array_equal(got["rgba"], ref["rgba"]) and np.array_equal(got["counters"], ref["counters"])


@pytest.mark.parametrize("seed", range(4))
def test_random_trees_through_the_whole_kernel_body(hc, seed):
    rng = np.random.default_rng(8800 + seed)
    W, H = 22, 14
    u = h.orbit_uniforms(W, H, yaw=0.4 * seed, pitch=0.3)
    for n_leaves, ext in ((3, False), (9, False), (4, True), (11, True)):
        words, cmd = h.random_program(rng, n_leaves, ext=ext)
        ref = h.oracle_render(u, cmd, words, W, H)
        got = paint(hc, cmd, words, u, W, H, counters=True)
        assert np.array_equal(got["rgba"], ref["rgba"]), (seed, n_leaves, ext)
        assert np.array_equal(got["counters"], ref["counters"])


def test_fast_mode_on_the_host_is_inside_the_gates(hc):
    # the host's fast mode replaces the MUFU approximations by correctly rounded operations, so this is a check of the fast
    # CODE PATH (FMA contraction, reciprocal constants, rsqrt normalisation), not of the hardware's approximations
    for name, W, H, yaw, pitch in CASES:
        cmd, words = h.SCENES[name]()
        u = h.orbit_uniforms(W, H, yaw=yaw, pitch=pitch)
        ref = h.oracle_render(u, cmd, words, W, H, aov=True)
        got = paint(hc, cmd, words, u, W, H, exact=False, aov=True)
        d = np.abs(got["rgba"].astype(int) - ref["rgba"].astype(int)).max(-1)
        assert (d > 2).mean() <= 2e-3, name
        assert ((got["cls"] == 2) != (ref["cls"] == 2)).any(-1).mean() <= 2e-3, name


def test_collapsed_lowering_through_the_whole_kernel_body_is_inside_the_gates(hc):
    # the fast form of extended programs (rm_program.h collapse_chains): one similarity per primitive
    cmd, words = h.scene_s3()
    W, H = 48, 27
    u = h.orbit_uniforms(W, H, yaw=0.6, pitch=0.4)
    ref = h.oracle_render(u, cmd, words, W, H, aov=True)
    got = paint(hc, cmd, words, u, W, H, exact=-1, aov=True)
    d = np.abs(got["rgba"].astype(int) - ref["rgba"].astype(int)).max(-1)
    assert (d > 2).mean() <= 2e-3
    assert ((got["cls"] == 2) != (ref["cls"] == 2)).any(-1).mean() <= 2e-3
    both = (got["cls"] == 2) & (ref["cls"] == 2)
    rel = np.abs(got["t"][both] - ref["t"][both]) / np.maximum(np.abs(ref["t"][both]), 1e-6)
    assert (rel > 1e-4).mean() <= 2e-3


# ---- the run-time generated kernel (rm_jit.cu): the emitted `rm_spec` function itself, compiled for the host -------------
SPEC_DRIVER = """
#include "rm_frame.h"
extern "C" int hc_spec_paint(const void *uniforms144, const void *limits12, uint32_t W, uint32_t H, uint32_t row0, uint32_t nrows,
                             uint32_t strip_rows, uint32_t strip_pitch, int dst_frame_rows, uint8_t *dst, size_t pitch,
                             uint64_t *counters4, uint8_t *aov_cls, float *aov_t) {
    rm::SpecParams P;
    memcpy(&P, hc_spec_params, sizeof P);
#ifdef RM_SPEC_CONSTBANK   /* structure tier of a deep program: the scene code reads the module's copy (launch_spec refills it) */
    memcpy(&g_rm_spec_P, hc_spec_params, sizeof g_rm_spec_P);
    memset(&P, 0xff, sizeof P);   /* the kernel argument is not read */
#endif
    float u[36], lim[2];
    uint32_t max_iter;
    memcpy(u, uniforms144, 144); memcpy(lim, limits12, 8); memcpy(&max_iter, (const char *)limits12 + 8, 4);
    rm::FrameParams fp = rm::make_frame_params(u, lim[0], lim[1], max_iter, 0, W, H, row0, nrows);
    fp.strip_rows = strip_rows; fp.strip_pitch = strip_pitch; fp.dst_frame_rows = dst_frame_rows ? 1u : 0u;
    const uint32_t tile_h = rm::direct_tile_rows(HC_R);
    hc_run_grid(rm::direct_grid(fp, tile_h), RM_DIRECT_THREADS, [&] { rm_spec(fp, P, dst, pitch, (unsigned long long *)counters4, aov_cls, aov_t); });
    return 0;
}
"""


def spec_paint(tmp_path, rm_lib, cmd, words, u, W, H, exact=True, R=0, literal=False, **kw):
    w = np.ascontiguousarray(np.asarray(words, np.uint32))
    buf = C.create_string_buffer(1 << 20)
    need = C.c_size_t()
    rc = rm_lib.rm_spec_source(cmd, w.ctypes.data_as(u32p), w.size, (0 if exact else 1) | (0x100 if literal else 0), R, buf, 1 << 20, C.byref(need))
    assert rc == 0, rc
    src = buf.value.decode()
    r = int(re.search(r"rm::SpecMap<(?:true|false), (\d)> map\(", src).group(1))
    cu = tmp_path / "spec_kernel.cpp"
    import test_hostcheck_cpu as thc
    params = thc._spec_params(rm_lib, cmd, w, exact)
    table = "static const float hc_spec_params[] = {" + ", ".join("%sf" % repr(float(x)) if np.isfinite(x) else "0.0f" for x in params) + "};\n"
    table = "static const unsigned hc_spec_params_bits[] = {" + ", ".join("0x%08xu" % int(x) for x in params.view(np.uint32)) + "};\n" \
            "#define hc_spec_params ((const float *)hc_spec_params_bits)\n"
    cu.write_text('#include "hc_grid.h"\n' + src + f"\n#define HC_R {r}\n" + table + SPEC_DRIVER)
    so = tmp_path / f"libspec_kernel_{abs(hash((cmd, bytes(w), exact, R, literal)))}.so"  # params are baked into the driver: one .so per word stream
    subprocess.run(CXX + ["-pthread", f"-I{HC_DIR}", "-o", str(so), str(cu)], check=True, capture_output=True)
    lib = C.CDLL(str(so))
    lib.hc_spec_paint.restype = C.c_int
    lib.hc_spec_paint.argtypes = PAINT_ARGS
    nrows = kw.pop("nrows", None)
    row0 = kw.pop("row0", 0)
    nrows = H - row0 if nrows is None else nrows
    return _call(lib.hc_spec_paint, (), u, W, H, row0, nrows, kw.get("strip_rows", 0), kw.get("strip_pitch", 0), False, H,
                 kw.get("limits"), False, False), r


@pytest.fixture(scope="module")
def rm_lib():
    so = h.ROOT / "ray-marching_b200" / "librm_b200.so"
    if not so.exists():
        pytest.skip("librm_b200.so not built")
    lib = C.CDLL(str(so))
    lib.rm_spec_source.restype = C.c_int
    lib.rm_spec_source.argtypes = [C.c_uint32, u32p, C.c_uint32, C.c_int, C.c_int, C.c_char_p, C.c_size_t, C.POINTER(C.c_size_t)]
    return lib


@pytest.mark.parametrize("name,W,H,yaw,pitch", CASES)
def test_generated_kernel_exact_mode_is_bit_identical_to_the_oracle(hc, rm_lib, tmp_path, name, W, H, yaw, pitch):
    cmd, words = h.SCENES[name]()
    u = h.orbit_uniforms(W, H, yaw=yaw, pitch=pitch)
    ref = h.oracle_render(u, cmd, words, W, H)
    got, r = spec_paint(tmp_path, rm_lib, cmd, words, u, W, H)
    assert r == (2 if name == "S2" else 4)          # the rays per thread the generator chose: all three variants get run
    assert np.array_equal(got["rgba"], ref["rgba"]), name
    # a band of the same frame
    got, _ = spec_paint(tmp_path, rm_lib, cmd, words, u, W, H, row0=5, nrows=9)
    assert np.array_equal(got["rgba"], ref["rgba"][5:14])


def test_generated_kernel_with_eight_rays_per_thread(hc, rm_lib, tmp_path):
    cmd, words = h.scene_s1()
    u = h.orbit_uniforms(19, 11, yaw=0.6, pitch=0.4)
    got, r = spec_paint(tmp_path, rm_lib, cmd, words, u, 19, 11, R=8)
    assert r == 8 and np.array_equal(got["rgba"], h.oracle_render(u, cmd, words, 19, 11)["rgba"])


@pytest.mark.parametrize("k", [0.05, 1.0, 4.0, 4.01, 9.0])
def test_fast_smooth_union_saturate_form_equals_the_max_form_bit_for_bit(hc, k):
    """Fast mode, 8 rays per thread: the common-tail interpreter clamps the blend term with a saturate when k*s <= 1
    (FLAG_SAT_OK, rm_program.h); 4 rays per thread: map_evalR always uses max(., 0).  Same frame, bit for bit, on both
    sides of k = 4 (beyond it neither uses the saturate)."""
    W, H = 64, 36
    u = h.orbit_uniforms(W, H, yaw=0.5, pitch=0.35, radius=7.0)
    words = h.sphere((-2.0, 0.0, 0.0), 0.7)
    for i in range(1, 10):
        leaf = h.box((-2.0 + 0.45 * i, 0.15 * (i % 3), 0.1 * i), (0.35, 0.3, 0.25)) if i % 2 else h.sphere((-2.0 + 0.45 * i, 0.2, 0.0), 0.45)
        words = words + leaf + [h.OP_SMOOTH_UNION, h.fbits(k)]
    a = paint(hc, 19, words, u, W, H, exact=0, R=8, aov=True)
    b = paint(hc, 19, words, u, W, H, exact=0, R=4, aov=True)
    assert np.array_equal(a["rgba"], b["rgba"]) and np.array_equal(a["cls"], b["cls"])
    assert np.array_equal(a["t"].view(np.uint32), b["t"].view(np.uint32))
    assert (a["cls"] == 2).any()      # the chain is in view


@pytest.mark.parametrize("name,W,H", [("S1", 40, 22), ("S2", 40, 22), ("chain", 40, 22)])
def test_generated_fast_kernel_in_ray_pairs_equals_the_interpreter_bit_for_bit(hc, rm_lib, tmp_path, name, W, H):
    """Fast mode: the generated scene code evaluates the rays of a thread in PAIRS (f2_t; on the GPU FADD2 / FMUL2 / FFMA2).
    Per lane it is the scalar fast-mode arithmetic, operation for operation, so the host build of the generated kernel —
    S1 inlined with 4 rays per thread, S2 out of line with 2 and the constant block, a smooth-union chain on both sides of
    k = 4 — gives the interpreter's fast-mode frame bit for bit."""
    if name == "chain":
        words = h.sphere((-2.0, 0.0, 0.0), 0.7)
        for i in range(1, 10):
            leaf = h.box((-2.0 + 0.45 * i, 0.15 * (i % 3), 0.1 * i), (0.35, 0.3, 0.25)) if i % 2 else h.sphere((-2.0 + 0.45 * i, 0.2, 0.0), 0.45)
            words = words + leaf + [h.OP_SMOOTH_UNION, h.fbits((0.3, 4.0, 9.0)[i % 3])]
        words = words + h.sphere((0.0, 0.4, 0.5), 0.5) + [h.OP_SUBTRACTION] + h.box((0, 0, 0), (3, 0.9, 0.9)) + [h.OP_INTERSECTION]
        cmd = 19 + 4
        u = h.orbit_uniforms(W, H, yaw=0.5, pitch=0.35, radius=7.0)
    else:
        cmd, words = h.SCENES[name]()
        u = h.orbit_uniforms(W, H, yaw=0.6, pitch=0.4)
    got, r = spec_paint(tmp_path, rm_lib, cmd, words, u, W, H, exact=False)
    ref = paint(hc, cmd, words, u, W, H, exact=0, R=8 if name != "S1" else 4)
    assert np.array_equal(got["rgba"], ref["rgba"]), (name, r)
    assert (got["rgba"][..., :3] != got["rgba"][0, 0, :3]).any()


@pytest.mark.parametrize("seed,n_leaves", [(0, 3), (1, 7), (2, 12), (3, 21)])
def test_generated_fast_kernel_in_ray_pairs_on_random_trees(hc, rm_lib, tmp_path, seed, n_leaves):
    """The same identity on random trees: every operator, smooth unions with several k, sub-trees on either side (spill
    slots in the interpreter, SSA values in the generated code)."""
    from test_gpu_parity import random_program
    rng = np.random.default_rng(900 + seed)
    words, cmd = random_program(rng, n_leaves)
    W, H = 26, 15
    u = h.orbit_uniforms(W, H, yaw=0.8 * seed, pitch=0.3, radius=5.0)
    got, r = spec_paint(tmp_path, rm_lib, cmd, words, u, W, H, exact=False)
    ref = paint(hc, cmd, words, u, W, H, exact=0)
    assert np.array_equal(got["rgba"], ref["rgba"]), (seed, n_leaves, r)


def test_generated_literal_tier_fast_kernel_equals_the_interpreter_bit_for_bit(hc, rm_lib, tmp_path):
    """The literal tier of the deep scene (every value an immediate, out of line, rays in pairs) on the host."""
    cmd, words = h.scene_s2()
    W, H = 40, 22
    u = h.orbit_uniforms(W, H, yaw=0.6, pitch=0.4)
    got, r = spec_paint(tmp_path, rm_lib, cmd, words, u, W, H, exact=False, literal=True)
    assert r == 2
    assert np.array_equal(got["rgba"], paint(hc, cmd, words, u, W, H, exact=0, R=8)["rgba"])
    got, _ = spec_paint(tmp_path, rm_lib, cmd, words, u, W, H, exact=True, literal=True)
    assert np.array_equal(got["rgba"], h.oracle_render(u, cmd, words, W, H)["rgba"])
