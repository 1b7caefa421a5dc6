"""rm_paint_culled_static — the allocation-free entry point (include/rm_b200.h; no counterpart in the reference):
the culled paint with its tile masks in a file-scope static __device__ array instead of a
cudaMalloc'ed buffer.  Called through the C ABI and compared with the existing entry points (rm_paint, with and without
rm_set_cull) on two calls in a row.  The paint is exact by construction, so the tolerance is the one the culling tests use:
byte identity.  Collected last (zz) so that a failure here hides no other test under `pytest -x`; passes alone in a fresh
process and after the other tests (the static array is shared, process-wide state: every call rewrites what it reads)."""
import numpy as np
import pytest

import helpers as h
import rm_import

pytestmark = pytest.mark.gpu

MAX_TILES = 262144  # RM_CULL_STATIC_MAX_TILES


@pytest.fixture(scope="module")
def rm():
    return rm_import.load()


def test_static_scratch_paint_equals_the_allocating_entry_points_twice_in_a_row(rm):
    import torch
    dev = torch.device("cuda", 0)
    cmd, words = h.scene_s2()
    # two calls in a row with DIFFERENT frames: a big one, then a smaller one that leaves most of the array holding the
    # first call's masks; then the first again
    frames = [(640, 360, 0.6, 0.4, 5.0), (200, 120, 2.4, 0.1, 3.0), (640, 360, 0.6, 0.4, 5.0)]
    with rm.Context(0) as ctx:
        for mode in (rm.MODE_EXACT, rm.MODE_FAST):
            ctx.set_mode(mode)
            refs = []
            for W, H, yaw, pitch, radius in frames:      # existing entry points: rm_paint, and rm_paint with rm_set_cull
                u = h.orbit_uniforms(W, H, yaw=yaw, pitch=pitch, radius=radius)
                ctx.set_cull(False)
                plain = ctx.render(u, cmd, words, W, H)
                ctx.set_cull(True)
                culled = ctx.render(u, cmd, words, W, H)
                ctx.set_cull(False)
                assert np.array_equal(plain, culled)
                refs.append((u, plain))
            bufs = [torch.full((H, W, 4), 7, dtype=torch.uint8, device=dev) for (W, H, *_r) in frames]   # the caller's outputs
            torch.cuda.synchronize(dev)
            # one call outside the measurement resolves the array's device address (first call per device)
            ctx.prepare(refs[0][0], cmd, words)
            ctx.paint_culled_static(frames[0][0], frames[0][1], 0, frames[0][1], bufs[0].data_ptr())
            ctx.sync()
            free0 = torch.cuda.mem_get_info(dev)[0]
            for (W, H, *_r), (u, ref), buf in zip(frames, refs, bufs):
                ctx.prepare(u, cmd, words)
                ctx.paint_culled_static(W, H, 0, H, buf.data_ptr())   # back to back on the ctx stream: serialised, no overlap
            ctx.sync()
            free1 = torch.cuda.mem_get_info(dev)[0]
            assert free1 >= free0, (free0, free1)                     # nothing was allocated on the device by the calls
            for (W, H, *_r), (u, ref), buf in zip(frames, refs, bufs):
                got = buf.cpu().numpy()
                assert np.array_equal(got, ref), (mode, W, H, int((got != ref).any(-1).sum()))
            if mode == rm.MODE_EXACT:                                  # and the oracle, for the first frame
                W, H = frames[0][:2]
                assert np.array_equal(bufs[0].cpu().numpy(), h.oracle_render(refs[0][0], cmd, words, W, H)["rgba"])
        # a band of a frame: rows 40..199 land at offset 0 of the caller's buffer, like rm_paint
        ctx.set_mode(rm.MODE_EXACT)
        W, H = frames[0][:2]
        ctx.prepare(refs[0][0], cmd, words)
        band = torch.zeros((160, W, 4), dtype=torch.uint8, device=dev)
        ctx.paint_culled_static(W, H, 40, 160, band.data_ptr())
        ctx.sync()
        assert np.array_equal(band.cpu().numpy(), ctx.render(refs[0][0], cmd, words, W, H)[40:200])


def test_static_scratch_limits_are_rejected_with_a_clear_error(rm):
    import torch
    dev = torch.device("cuda", 0)
    cmd, words = h.scene_s2()
    with rm.Context(0) as ctx:
        ctx.set_mode(rm.MODE_FAST)
        # 7680 x 4320 = 960 x 540 = 518400 tiles of 8 x 8 pixels > 262144
        W, H = 7680, 4320
        ctx.prepare(h.orbit_uniforms(W, H, yaw=0.6, pitch=0.4), cmd, words)
        small = torch.zeros(64, dtype=torch.uint8, device=dev)   # never written: the call must fail before any launch
        with pytest.raises(rm.RmError) as e:
            ctx.paint_culled_static(W, H, 0, H, small.data_ptr())
        assert e.value.code == -1 and "518400" in str(e.value) and str(MAX_TILES) in str(e.value) and "row bands" in str(e.value)
        # the same frame in two row bands of 2160 rows each fits (259200 tiles per band): check the upper band's rows
        band = torch.zeros((2160, W, 4), dtype=torch.uint8, device=dev)
        ctx.paint_culled_static(W, H, 0, 2160, band.data_ptr())
        ctx.sync()
        rows = [0, 777, 2159]
        ref = np.stack([ctx.render(h.orbit_uniforms(W, H, yaw=0.6, pitch=0.4), cmd, words, W, H, row0=r, nrows=1)[0] for r in rows])
        assert np.array_equal(band[torch.as_tensor(rows, device=dev)].cpu().numpy(), ref)
        # no buffer of the caller's: refused (this entry point allocates nothing)
        with pytest.raises(rm.RmError) as e:
            ctx.paint_culled_static(64, 36, 0, 36, None)
        assert e.value.code == -1 and "allocates nothing" in str(e.value)
        # programs the culled paint does not cover
        c1, w1 = h.scene_s1()
        ctx.prepare(h.orbit_uniforms(64, 36), c1, w1)
        out = torch.zeros((36, 64, 4), dtype=torch.uint8, device=dev)
        with pytest.raises(rm.RmError) as e:
            ctx.paint_culled_static(64, 36, 0, 36, out.data_ptr())
        assert e.value.code == rm.E_UNSUPPORTED
        c3, w3 = h.scene_s3()
        ctx.prepare(h.orbit_uniforms(64, 36), c3, w3)
        with pytest.raises(rm.RmError) as e:
            ctx.paint_culled_static(64, 36, 0, 36, out.data_ptr())
        assert e.value.code == rm.E_UNSUPPORTED
