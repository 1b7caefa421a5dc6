#!/usr/bin/env python
"""A small tour of every kernel and copy path of the library through the C ABI (no torch) in ONE process: ragged frames of
a few thousand pixels, every kernel generation, both arithmetic modes, both specialised tiers, the culled paint and its
allocation-free twin, bands, strips (packed and into a whole frame), pose batches to host memory, AOVs, counters and
rm_map, every result compared with the interpreter's frame.  Written for compute-sanitizer (memcheck / racecheck /
initcheck); that tool is closed on the GPU pool of this build (`profiles/r2_s33_sanitizer_*.log`), so the bounds check of
the kernel body's stores runs on the host instead (tests/test_hostcheck_asan_cpu.py: the same body under AddressSanitizer)
and this tour runs plain.

  python tools/sanitize_driver.py                                  # must print "tour ok"
"""
import importlib
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import rm_import  # noqa: E402


def main():
    rm = rm_import.load()
    scenes = importlib.import_module("ray_marching_b200.scenes")
    sh = importlib.import_module("ray_marching_b200.sharding")
    W, H = 76, 43                      # ragged: not a multiple of any tile (W a multiple of 4: rm_paint_strips_into wants a 16-byte pitch)
    n_paints = 0
    with rm.Context(0) as ctx:
        for scene in ("S0", "S1", "S2", "S3"):
            cmd, words = scenes.program(scene)
            u = scenes.orbit_uniforms(W, H, yaw=0.6, pitch=0.4)
            for mode in (rm.MODE_EXACT, rm.MODE_FAST):
                ctx.set_mode(mode)
                ctx.set_kernel(rm.KERNEL_DIRECT)
                ctx.set_specialize(False)
                ref = ctx.render(u, cmd, words, W, H)
                n_paints += 1
                # the other kernel generations (exact mode: identical images)
                for k in (rm.KERNEL_WAVEFRONT2, rm.KERNEL_WAVEFRONT, rm.KERNEL_SIMPLE):
                    ctx.set_kernel(k)
                    try:
                        img = ctx.render(u, cmd, words, W, H)
                    except rm.RmError as e:   # a kernel generation that does not take this program (extended commands)
                        assert e.code == rm.E_UNSUPPORTED, e
                        continue
                    n_paints += 1
                    if mode == rm.MODE_EXACT:
                        assert np.array_equal(img, ref), (scene, "kernel", k)
                ctx.set_kernel(rm.KERNEL_DIRECT)
                # both specialised tiers
                for after in (0, 1):
                    ctx.set_spec_literal_after(after)
                    ctx.prepare(u, cmd, words)
                    ctx.specialize()
                    for _ in range(2):
                        ctx.paint(W, H)
                        n_paints += 1
                    img = ctx.download(np.zeros((H, W, 4), np.uint8))
                    if mode == rm.MODE_EXACT:
                        assert np.array_equal(img, ref), (scene, "tier", ctx.spec_tier)
                    a = ctx.paint_aov(W, H)
                    n_paints += 1
                    assert np.array_equal(a["rgba"], img)
                ctx.set_spec_literal_after(8)
                ctx.set_specialize(False)
                # counters, AOVs, rm_map through the interpreter
                ctx.set_counters(True)
                ctx.render(u, cmd, words, W, H)
                assert int(ctx.get_counters()[3]) == W * H * 16
                ctx.set_counters(False)
                ctx.paint_aov(W, H)
                ctx.paint_debug(W, H)
                ctx.map(np.random.default_rng(1).uniform(-3, 3, (257, 3)).astype(np.float32))
                n_paints += 3
                # a band, packed strips, strips into a whole frame
                assert np.array_equal(ctx.render(u, cmd, words, W, H, row0=7, nrows=19), ref[7:26])
                world, strip = 3, 5
                gathered = np.zeros((world, sh.max_local_rows(H, strip, world), W, 4), np.uint8)
                ctx.prepare(u, cmd, words)
                for r in range(world):
                    n = ctx.paint_strips(W, H, strip, r, world)
                    ctx.download(gathered[r, :n])
                    n_paints += 1
                assert np.array_equal(sh.reassemble(gathered, H, strip, world), ref)
                ptr, _handle = ctx.shared_frame_create(W * H * 4)
                for r in range(world):
                    ctx.paint_strips_into(W, H, strip, r, world, ptr, W * 4)
                    n_paints += 1
                ctx.sync()
                frame = np.zeros((H, W, 4), np.uint8)
                ctx.copy_to_host(ptr, frame)
                assert np.array_equal(frame, ref), (scene, "strips_into")
                host = rm.PinnedBuffer(W * H * 4)
                for r in range(world):
                    ctx.paint_strips(W, H, strip, r, world)
                    ctx.copy_strips_to_host(None, W, H, strip, r, world, host.ptr)
                assert np.array_equal(host.array.reshape(H, W, 4), ref), (scene, "copy_strips_to_host")
                host.free()
                # culled paint and the allocation-free entry point (programs of 8..64 reference-set instructions only)
                ctx.set_cull(True)
                ctx.cull_stats()
                img = ctx.render(u, cmd, words, W, H)
                n_paints += 1
                assert np.array_equal(img, ref), (scene, "cull")
                if ctx.cull_stats()["culled_paints"]:
                    ctx.prepare(u, cmd, words)
                    ctx.paint_culled_static(W, H, 0, H, ptr)
                    ctx.sync()
                    ctx.copy_to_host(ptr, frame)
                    assert np.array_equal(frame, ref), (scene, "culled_static")
                    n_paints += 1
                ctx.set_cull(False)
                ctx.shared_frame_close(ptr)
                # a pose batch to host memory (the pipelined path)
                poses = [scenes.orbit_uniforms(W, H, yaw=0.3 * k, pitch=0.4) for k in range(5)]
                pb = rm.PinnedBuffer(len(poses) * W * H * 4)
                ctx.prepare(poses[0], cmd, words)
                ctx.paint_poses_host(b"".join(poses), len(poses), W, H, pb.ptr)
                got = pb.array.reshape(len(poses), H, W, 4).copy()
                pb.free()
                n_paints += len(poses)
                for k, pu in enumerate(poses):
                    assert np.array_equal(ctx.render(pu, cmd, words, W, H), got[k]), (scene, "pose", k)
        # limits and an empty program
        ctx.set_limits(0.05, 40.0, 7)
        ctx.render(scenes.orbit_uniforms(W, H), 0, [], W, H)
        ctx.set_limits(0.01, 100.0, 100)
    print(f"tour ok: {n_paints} paints of {W}x{H}")
    return 0


if __name__ == "__main__":
    sys.exit(main())
