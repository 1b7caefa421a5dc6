"""rm_group_*: ONE host thread drives several GPUs (SURVEY.md 8(b) "Multi-GPU (new surface)", 8(e)).
On the driver's one-GPU box the members are several contexts on cuda:0 (listing a device twice is allowed): the strip
plan, the frame-row stores into the first member's frame, the round-robin pose split and the host-slot addressing are
the real ones; with >= 2 GPUs the members sit on different devices and the stores cross NVLink.
Everything must be bit-identical to single-context renders."""
import numpy as np
import pytest
import torch

import helpers as h
import rm_import

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def rm():
    return rm_import.load()


def _devices(n):
    have = max(1, torch.cuda.device_count())
    return [k % have for k in range(n)]


@pytest.mark.parametrize("n", [1, 2, 3])
@pytest.mark.parametrize("W,H,strip", [(256, 150, 16), (200, 77, 4), (64, 40, 64)])
def test_group_bands_equal_the_single_context_frame(rm, n, W, H, strip):
    cmd, words = h.scene_s2()
    u = h.orbit_uniforms(W, H, yaw=0.6, pitch=0.4)
    with rm.Context(0) as ctx:
        ctx.set_mode(rm.MODE_FAST)
        full = ctx.render(u, cmd, words, W, H)
    with rm.Group(_devices(n)) as g:
        assert len(g) == n
        g.set_mode(rm.MODE_FAST)
        g.prepare(u, cmd, words)
        got = np.zeros((H, W, 4), np.uint8)
        g.paint_bands(W, H, strip, got)
        assert np.array_equal(got, full)
        # the frame is also complete on the first device, pitch rounded up to 16 bytes
        ptr, pitch = g.frame()
        assert ptr and pitch >= W * 4 and pitch % 16 == 0
        # every member with at least one strip launched work
        n_strips = (H + strip - 1) // strip
        for i in range(n):
            assert (g.member_launch_count(i) > 0) == (i < n_strips)
        # a second frame with another pose reuses the uploaded program
        u2 = h.orbit_uniforms(W, H, yaw=2.0, pitch=0.2)
        g.prepare(u2, cmd, words)
        got2 = np.zeros((H, W, 4), np.uint8)
        g.paint_bands(W, H, strip, got2)
    with rm.Context(0) as ctx:
        ctx.set_mode(rm.MODE_FAST)
        assert np.array_equal(got2, ctx.render(u2, cmd, words, W, H))


def test_group_bands_exact_mode_matches_the_oracle(rm):
    W, H = 96, 64
    cmd, words = h.scene_s1()
    u = h.orbit_uniforms(W, H, yaw=0.3, pitch=0.2)
    with rm.Group(_devices(2)) as g:
        g.set_mode(rm.MODE_EXACT)
        g.set_limits(0.01, 100.0, 100)
        g.prepare(u, cmd, words)
        got = np.zeros((H, W, 4), np.uint8)
        g.paint_bands(W, H, 8, got)
    assert np.array_equal(got, h.oracle_render(u, cmd, words, W, H)["rgba"])


@pytest.mark.parametrize("n,n_poses", [(1, 5), (2, 7), (3, 8), (4, 2)])
def test_group_pose_batch_lands_in_pose_order(rm, n, n_poses):
    W, H = 160, 90
    cmd, words = h.scene_s2()
    us = [h.orbit_uniforms(W, H, yaw=0.9 * k, pitch=0.4) for k in range(n_poses)]
    with rm.Context(0) as ctx:
        ctx.set_mode(rm.MODE_FAST)
        singles = [ctx.render(p, cmd, words, W, H) for p in us]
    buf = rm.PinnedBuffer(n_poses * W * H * 4)
    buf.array[:] = 0
    with rm.Group(_devices(n)) as g:
        g.set_mode(rm.MODE_FAST)
        g.prepare(us[0], cmd, words)
        g.paint_poses(b"".join(us), n_poses, W, H, buf.ptr)
    frames = buf.array.reshape(n_poses, H, W, 4)
    for k in range(n_poses):
        assert np.array_equal(frames[k], singles[k]), k
    buf.free()


def test_group_rejects_bad_arguments(rm):
    with pytest.raises(rm.RmError):
        rm.Group([])
    with pytest.raises(rm.RmError):
        rm.Group([torch.cuda.device_count() + 7])
    with rm.Group([0]) as g:
        cmd, words = h.scene_s0()
        g.prepare(h.orbit_uniforms(32, 32), cmd, words)
        with pytest.raises(rm.RmError):
            g.paint_bands(32, 32, 0)
        with pytest.raises(rm.RmError):
            g.prepare(h.orbit_uniforms(32, 32), 1, [5])  # unknown opcode: the member's message comes through
        with pytest.raises(rm.RmError):
            g.paint_poses(h.orbit_uniforms(32, 32), 1, 32, 32, None)
