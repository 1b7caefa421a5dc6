"""N>1 host logic on CPU: strip plans, and a world_size-2 gloo run in which each rank produces its
packed strips (with the CPU oracle standing in for the device), rank 0 gathers and reassembles, and
the result must equal the full frame bit for bit."""
import importlib
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import helpers as h
import rm_import


def sharding():
    rm_import.load()
    return importlib.import_module("ray_marching_b200.sharding")


def test_balanced_strip_height_gives_every_rank_the_same_number_of_strips():
    sh = sharding()
    assert sh.balanced_strip_rows(4320, 1) == 16 and sh.balanced_strip_rows(4320, 2) == 16      # 270 strips: 135 each
    assert sh.balanced_strip_rows(4320, 8) == 12 and sh.balanced_strip_rows(4320, 4) == 12      # 360 strips: 45 / 90 each
    assert sh.balanced_strip_rows(2160, 8) == 10                                                 # 216 strips: 27 each
    assert sh.balanced_strip_rows(1080, 7) == 16                                                 # nothing balances: the default
    for H, world in ((4320, 8), (4320, 4), (4320, 2), (2160, 8), (2160, 3), (1080, 5)):
        s = sh.balanced_strip_rows(H, world)
        rows = [sh.local_rows(H, s, r, world) for r in range(world)]
        assert sum(rows) == H and 4 <= s <= 16
        if H % s == 0 and (H // s) % world == 0:
            assert len(set(rows)) == 1, (H, world, s, rows)


@pytest.mark.parametrize("H,strip,world", [(4320, 16, 8), (4320, 12, 8), (2160, 15, 8), (2160, 16, 4), (100, 16, 3), (7, 4, 2), (33, 8, 8), (16, 16, 2)])
def test_strip_plan_covers_every_row_once(H, strip, world):
    sh = sharding()
    assert sum(sh.local_rows(H, strip, r, world) for r in range(world)) == H
    rank_idx, loc_idx = sh.row_index(H, strip, world)
    seen = set(zip(rank_idx.tolist(), loc_idx.tolist()))
    assert len(seen) == H
    for r in range(world):
        loc = sorted(l for (rr, l) in seen if rr == r)
        assert loc == list(range(sh.local_rows(H, strip, r, world)))  # packed, no holes
    assert list(sh.pose_slice(10, 1, 4)) == [1, 5, 9]


def _worker(rank, world, port, W, H, strip, out_path):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    sh = sharding()
    cmd, words = h.scene_s1()
    u = h.orbit_uniforms(W, H, yaw=0.6, pitch=0.4)
    rows_max = sh.max_local_rows(H, strip, world)
    band = torch.zeros((rows_max, W, 4), dtype=torch.uint8)
    r0 = 0
    for s in range(rank, sh.n_strips(H, strip), world):  # what rm_paint_strips does on the device
        n = min(strip, H - s * strip)
        band[r0:r0 + n] = torch.from_numpy(h.oracle_render(u, cmd, words, W, H, row0=s * strip, nrows=n, nthreads=1)["rgba"])
        r0 += n
    assert r0 == sh.local_rows(H, strip, rank, world)
    gathered = [torch.zeros_like(band) for _ in range(world)] if rank == 0 else None
    dist.gather(band, gathered, dst=0)
    if rank == 0:
        frame = sh.reassemble(torch.stack(gathered), H, strip, world).numpy()
        np.save(out_path, frame)
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_gloo_gather_reassembles_the_frame(tmp_path):
    W, H, strip, world = 64, 43, 4, 2
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    out = str(tmp_path / "frame.npy")
    mp.spawn(_worker, args=(world, port, W, H, strip, out), nprocs=world, join=True)
    cmd, words = h.scene_s1()
    ref = h.oracle_render(h.orbit_uniforms(W, H, yaw=0.6, pitch=0.4), cmd, words, W, H)["rgba"]
    assert np.array_equal(np.load(out), ref)
