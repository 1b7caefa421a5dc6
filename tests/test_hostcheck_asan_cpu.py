"""The production kernel body under AddressSanitizer on the host (tests/hostcheck/asan_main.cpp).  compute-sanitizer is
closed on the GPU pool, so this is the bounds check of the body's index arithmetic: ragged frames, bands, packed strips,
strips stored into a whole frame, counters and AOV arrays — each output a malloc of exactly the documented size."""
import struct
import subprocess
from pathlib import Path

import numpy as np
import pytest

import helpers as h

HC_DIR = Path(__file__).resolve().parent / "hostcheck"
CSRC = h.ROOT / "ray-marching_b200" / "csrc"


def build_exe():
    out = HC_DIR / "_build" / "hostcheck_asan"
    out.parent.mkdir(exist_ok=True)
    deps = [HC_DIR / "asan_main.cpp", HC_DIR / "hostcheck.cpp", HC_DIR / "hc_grid.h", HC_DIR / "shim" / "cuda_runtime.h"] + [
        CSRC / f for f in ("rm_program.cpp", "rm_program.h", "rm_device.cuh", "rm_wavefront_common.cuh", "rm_wavefront.h",
                           "rm_direct_body.cuh", "rm_frame.h")]
    if not out.exists() or out.stat().st_mtime < max(d.stat().st_mtime for d in deps):
        subprocess.run(["g++", "-O1", "-g", "-std=c++17", "-ffp-contract=off", "-pthread", "-fsanitize=address", "-fno-omit-frame-pointer",
                        f"-I{HC_DIR / 'shim'}", f"-I{CSRC}", "-w", "-o", str(out), str(HC_DIR / "asan_main.cpp"),
                        str(HC_DIR / "hostcheck.cpp")], check=True, capture_output=True)
    return out


def case(exact, rays, cmd, words, u, W, H, row0, nrows, strip_rows=0, strip_pitch=0, into_frame=False, variant=0, limits=None,
         dst_rows=None):
    words = np.asarray(words, np.uint32)
    dst_rows = (H if into_frame else nrows) if dst_rows is None else dst_rows
    head = struct.pack("<iiIIIIIIIIiIi", exact, rays, cmd, len(words), W, H, row0, nrows, strip_rows, strip_pitch,
                       1 if into_frame else 0, dst_rows, variant)
    return head + words.tobytes() + bytes(u) + (limits or h.limits_bytes(0.01, 100.0, 100))


def local_rows(H, strip, first, step):
    return sum(min(strip, H - s * strip) for s in range(first, (H + strip - 1) // strip, step))


def test_kernel_body_stores_stay_inside_exactly_sized_buffers(tmp_path):
    exe = build_exe()
    cases = []
    for name, W, H in (("S1", 37, 23), ("S2", 29, 18), ("S3", 21, 13), ("S0", 9, 5)):
        cmd, words = h.SCENES[name]()
        u = h.orbit_uniforms(W, H, yaw=0.6, pitch=0.4)
        for exact in (1, 0):
            for rays in (4, 8):
                cases.append(case(exact, rays, cmd, words, u, W, H, 0, H))                       # whole ragged frame
        cases.append(case(1, 0, cmd, words, u, W, H, 3, H - 5))                                  # a band
        cases.append(case(0, 0, cmd, words, u, W, H, 0, H, variant=1))                           # counters
        cases.append(case(1, 0, cmd, words, u, W, H, 2, H - 2, variant=2))                       # AOV arrays of a band
        for strip, step in ((4, 3), (5, 2), (16, 8)):
            for first in range(min(step, (H + strip - 1) // strip)):
                n = local_rows(H, strip, first, step)
                cases.append(case(0, 0, cmd, words, u, W, H, first * strip, n, strip, step * strip))                   # packed strips
                cases.append(case(1, 0, cmd, words, u, W, H, first * strip, n, strip, step * strip, into_frame=True))  # into a frame
    cases.append(case(1, 0, 0, [], h.orbit_uniforms(8, 8), 8, 8, 0, 8))                           # empty program
    cases.append(case(0, 4, *h.scene_s1(), h.orbit_uniforms(130, 3), 130, 3, 0, 3, limits=h.limits_bytes(0.05, 40.0, 7)))
    path = tmp_path / "cases.bin"
    path.write_bytes(struct.pack("<I", len(cases)) + b"".join(cases))
    r = subprocess.run([str(exe), str(path)], capture_output=True, text=True, timeout=1500)
    assert "AddressSanitizer" not in r.stderr, r.stderr[-3000:]
    assert r.returncode == 0, (r.returncode, r.stderr[-2000:])
    assert f"asan tour ok: {len(cases)} cases" in r.stdout


def test_the_check_has_teeth_a_buffer_one_row_short_is_reported(tmp_path):
    exe = build_exe()
    cmd, words = h.scene_s1()
    W, H = 24, 11
    path = tmp_path / "bad.bin"
    path.write_bytes(struct.pack("<I", 1) + case(0, 0, cmd, words, h.orbit_uniforms(W, H), W, H, 0, H, dst_rows=H - 1))
    r = subprocess.run([str(exe), str(path)], capture_output=True, text=True, timeout=600)
    assert r.returncode != 0 and "heap-buffer-overflow" in r.stderr
