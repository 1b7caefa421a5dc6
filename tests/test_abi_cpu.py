"""CPU-side checks of the C-ABI library: it loads, exports every declared symbol, and its static
program validation agrees with the oracle's.  No compute calls (no GPU here)."""
import ctypes as C
import re
from pathlib import Path

import numpy as np
import pytest

import helpers as h
import rm_import

ROOT = Path(__file__).resolve().parent.parent


@pytest.fixture(scope="module")
def rm():
    m = rm_import.load()
    if not m.LIB_PATH.exists():
        m.build()
    return m


def test_library_exports_every_declared_symbol(rm):
    header = (ROOT / "include" / "rm_b200.h").read_text()
    declared = set(re.findall(r"RM_API\s+[\w \*]+?\b(rm_\w+)\s*\(", header))
    assert declared == set(rm.ABI_SYMBOLS), declared ^ set(rm.ABI_SYMBOLS)
    L = rm.lib()
    for s in declared:
        assert hasattr(L, s), s
    assert b"sm_100a" in L.rm_version()


def test_strerror_covers_codes(rm):
    L = rm.lib()
    for code in range(0, -10, -1):
        assert L.rm_strerror(code) != b"unknown error"


@pytest.mark.parametrize("words,count,code", [
    ([], 0, 0),
    (h.sphere((0, 0, 0), 1), 1, 0),
    ([5], 1, -3),
    ([100], 1, -5),
    ([0, 0, 0, 0], 1, -5),
    (h.sphere((0, 0, 0), 1) * 33, 33, -4),
    (h.sphere((0, 0, 0), 1) * 32 + [100] * 31, 63, 0),
    (h.sphere((0, 0, 0), 1) * 60, 60, -2),
    (h.sphere((0, 0, 0), 1) * 2 + [103, h.fbits(0.0)], 3, -8),
    (h.sphere((0, 0, 0), 1) * 2 + [103, h.fbits(float("nan"))], 3, -8),
    (h.sphere((0, 0, 0), 1) * 2 + [103, h.fbits(0.25)], 3, 0),
    # NaN parameters are refused (the shader's result would be NaN everywhere); infinite radii are accepted (a slab)
    ([0, h.fbits(float("nan")), 0, 0, h.fbits(1.0)], 1, -8),
    ([1, 0, 0, 0, h.fbits(1.0), h.fbits(float("nan")), h.fbits(1.0)], 1, -8),
    ([1, 0, 0, 0, h.fbits(float("inf")), h.fbits(1.0), h.fbits(float("inf"))], 1, 0),
    ([200, 0, h.fbits(float("nan")), 0] + h.sphere((0, 0, 0), 1) + [201], 3, -8),
    # N4 (plane, space-transform groups)
    (h.plane((0, 1, 0), 1.5), 1, 0),
    (h.translated((1, 2, 3), (h.sphere((0, 0, 0), 1), 1))[0], 3, 0),
    (h.scaled(2.0, h.rotated(np.eye(3), (h.box((0, 0, 0), (1, 1, 1)), 1)))[0], 5, 0),
    ([201], 1, -5),
    ([200, 0, 0, 0] + h.sphere((0, 0, 0), 1) + [203], 3, -5),
    ([204, h.fbits(2.0)] + h.sphere((0, 0, 0), 1) + [205, h.fbits(4.0)], 3, -8),
    ([204, h.fbits(-2.0)] + h.sphere((0, 0, 0), 1) + [205, h.fbits(-2.0)], 3, -8),
    ([200, 0, 0, 0] * 9 + h.sphere((0, 0, 0), 1) + [201] * 9, 19, -4),
])
def test_validate_program_codes(rm, words, count, code):
    assert rm.validate_program(count, words) == code


def test_nan_and_bad_k_are_refused_by_the_oracle_validator_too():
    assert h.oracle_validate(1, [0, h.fbits(float("nan")), 0, 0, h.fbits(1.0)]) != 0
    assert h.oracle_validate(1, [1, 0, 0, 0, h.fbits(float("inf")), h.fbits(1.0), h.fbits(float("inf"))]) == 0
    two = h.sphere((0, 0, 0), 1) * 2
    for k, ok in ((0.25, True), (0.0, False), (-1.0, False), (float("inf"), False), (float("nan"), False), (3.2e38, True)):
        assert (h.oracle_validate(3, two + [103, h.fbits(k)]) == 0) == ok, k


def test_validation_agrees_with_oracle_on_random_streams(rm):
    rng = np.random.default_rng(0)
    ops = [0, 1, 100, 101, 102, 103, 7, 2, 200, 201, 202, 203, 204, 205]
    agree = 0
    for it in range(6000):
        n = int(rng.integers(1, 12))
        words = []
        for _ in range(n):
            t = ops[rng.integers(7 if it < 3000 else len(ops))]
            words.append(t)
            val = [0.5, 0.25][rng.integers(2)] if it >= 3000 else rng.uniform(0.1, 1)
            if rng.random() < 0.03:
                val = [float("nan"), float("inf"), 0.0, -0.5, 3.2e38][rng.integers(5)]
            words += [h.fbits(val)] * {
                0: 4, 1: 6, 103: 1, 2: 4, 200: 3, 202: 9, 204: 1, 205: 1}.get(t, 0)
        if rng.random() < 0.1:
            words = words[:-1]
        a = rm.validate_program(n, words) == 0
        b = h.oracle_validate(n, words) == 0
        assert a == b, words
        agree += a
    assert agree > 50


def test_scene_definitions():
    c, w = h.scene_s1()
    assert c == 5 and len(w) == 19
    c, w = h.scene_s2()
    assert c == 63 and len(w) == 251 and rm_import.load().validate_program(c, w) == 0


def test_group_entry_points_reject_null_arguments_without_a_device(rm):
    L = rm.lib()
    hnd = C.c_void_p()
    assert L.rm_group_create(None, 2, C.byref(hnd)) == -1
    assert L.rm_group_create((C.c_int * 1)(0), 0, C.byref(hnd)) == -1
    assert L.rm_group_size(None) == 0 and not L.rm_group_ctx(None, 0)
    assert L.rm_group_prepare(None, None, 0, None, 0) == -1
    assert L.rm_group_paint_bands(None, 8, 8, 4, None) == -1
    assert L.rm_group_paint_poses(None, None, 0, 8, 8, None, 0) == -1
    assert L.rm_group_last_error(None) == b"null group"
    L.rm_group_destroy(None)


def _arg_count(arglist: str) -> int:
    arglist = arglist.strip()
    return 0 if arglist in ("", "void") else arglist.count(",") + 1


def test_rust_binding_declares_the_same_functions_with_the_same_arity(rm):
    """bindings/rust/rm_sys.rs cannot be compiled here (no Rust toolchain): keep it in step with the header by name
    and by number of arguments."""
    header = (ROOT / "include" / "rm_b200.h").read_text()
    header = re.sub(r"/\*.*?\*/", "", header, flags=re.S)
    c_decl = {m.group(1): _arg_count(m.group(2))
              for m in re.finditer(r"RM_API\s+[\w \*]+?\b(rm_\w+)\s*\(([^)]*)\)\s*;", header)}
    rust = (ROOT / "bindings" / "rust" / "rm_sys.rs").read_text()
    rust = re.sub(r"//.*", "", rust)
    r_decl = {m.group(1): _arg_count(m.group(2)) for m in re.finditer(r"pub fn (rm_\w+)\s*\(([^)]*)\)", rust)}
    assert set(c_decl) == set(rm.ABI_SYMBOLS)
    assert r_decl == c_decl, {k: (c_decl.get(k), r_decl.get(k)) for k in set(c_decl) | set(r_decl) if c_decl.get(k) != r_decl.get(k)}


def test_python_constants_agree_with_the_header(rm):
    hdr = (Path(__file__).resolve().parent.parent / "include" / "rm_b200.h").read_text()
    seen = 0
    for name, val in re.findall(r"RM_(OP_\w+|E_\w+|MODE_\w+|KERNEL_\w+) = (-?\d+)", hdr):
        if hasattr(rm, name):
            assert getattr(rm, name) == int(val), (name, val)
            seen += 1
    assert seen >= 20
