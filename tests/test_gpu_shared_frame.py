"""rm_paint_strips_into + rm_shared_frame_*: every rank stores its strips straight into rank 0's frame (CUDA IPC mapping,
NVLink between GPUs) — no gather, no re-order.  The driver's GPU box has one GPU, so the two ranks here are two PROCESSES on
cuda:0 exchanging the 64-byte handle over gloo: the cross-process mapping and the frame-row addressing are the real ones;
with >= 2 GPUs the second rank uses cuda:1 and the stores cross the link.  Result: bit-identical to a full-frame render."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import helpers as h
import rm_import

pytestmark = pytest.mark.gpu


def _worker(rank, world, port, W, H, strip, kernel, out_path):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    rm = rm_import.load()
    dev = rank % max(1, torch.cuda.device_count())
    ctx = rm.Context(dev)
    ctx.set_mode(rm.MODE_FAST)
    ctx.set_kernel(getattr(rm, "KERNEL_" + kernel))
    cmd, words = h.scene_s2()
    u = h.orbit_uniforms(W, H, yaw=0.6, pitch=0.4)
    ctx.prepare(u, cmd, words)
    box = [None]
    if rank == 0:
        ptr, box[0] = ctx.shared_frame_create(W * H * 4)
    dist.broadcast_object_list(box, src=0)
    if rank != 0:
        ptr = ctx.shared_frame_open(box[0])
    sh = __import__("importlib").import_module("ray_marching_b200.sharding")
    n = ctx.paint_strips_into(W, H, strip, rank, world, ptr)
    assert n == sh.local_rows(H, strip, rank, world)
    ctx.sync()       # this rank's stores are complete ...
    dist.barrier()   # ... and so are everybody else's
    if rank == 0:
        got = np.empty((H, W, 4), np.uint8)
        ctx.copy_to_host(ptr, got)
        full = ctx.render(u, cmd, words, W, H)
        np.save(out_path, np.array([int(np.array_equal(got, full)), int((got != full).any(-1).sum())]))
    dist.barrier()
    if rank != 0:
        ctx.shared_frame_close(ptr)
    ctx.close()
    dist.destroy_process_group()


@pytest.mark.parametrize("kernel", ["DIRECT", "WAVEFRONT2", "SIMPLE"])
@pytest.mark.parametrize("W,H,strip,world", [(256, 150, 16, 2), (200, 77, 4, 3)])
def test_strips_stored_into_a_shared_frame_equal_the_full_render(tmp_path, W, H, strip, world, kernel):
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    out = str(tmp_path / "ok.npy")
    mp.spawn(_worker, args=(world, port, W, H, strip, kernel, out), nprocs=world, join=True)
    ok, n_bad = np.load(out)
    assert ok == 1, f"{n_bad} pixels differ"
