"""GPU: the reference-shaped host path (RayMarchingResources + RayMarchingCallback::prepare/paint in
the C++ host) must give the same frame as the raw C ABI and as the CPU oracle on the same bytes."""
import importlib

import numpy as np
import pytest

import helpers as h
import rm_import

pytestmark = pytest.mark.gpu


def test_callback_prepare_paint_matches_oracle():
    rm = rm_import.load()
    host = importlib.import_module("ray_marching_b200.host")
    W, H = 160, 120
    node = host.scene_s1()
    orbit = host.Orbit((0, 0, 0), 5.0)
    orbit.set(0.6, 0.4)
    pos, q = orbit.camera()
    res = host.lib().rmh_resources_new(0)
    assert res >= 0, host.lib().rmh_last_error()
    # the callback keeps the ctx in its default RM_MODE_EXACT
    img = host.callback_render(res, node, pos, q, W, H)
    cmd, words = host.build_commands(node)
    u = orbit.uniforms(W, H)
    ref = h.oracle_render(u, cmd, words, W, H)["rgba"]
    assert np.array_equal(img, ref)
    # empty scene: Option<CSGNode>::None (renderer.rs:225) -> floor + sky only
    img0 = host.callback_render(res, host.NONE, pos, q, W, H)
    assert np.array_equal(img0, h.oracle_render(u, 0, [], W, H)["rgba"])
    # the same ctx through the raw ABI
    with rm.Context(0) as ctx:
        assert np.array_equal(ctx.render(u, cmd, words, W, H), ref)
    host.lib().rmh_reset()
