"""CPU: the multi-GPU host logic with world_size 2 over gloo -- sample split, framebuffer reduce, and that the combined
image equals the single-process image of the same samples.  The per-rank "renderer" here is the CPU oracle (this is a
test of the host-side partition/reduce logic, not of the kernels)."""
import os
import socket
import sys

import numpy as np
import pytest

from conftest import ROOT
from gpupathtracer_b200 import multigpu as M


def test_sample_range_partition():
    for spp in (0, 1, 2, 7, 64, 1024, 1000):
        for world in (1, 2, 3, 4, 8):
            got = [M.sample_range(r, world, spp) for r in range(world)]
            assert got[0][0] == 0
            for (a, n), (b, _) in zip(got, got[1:]):
                assert a + n == b
            assert got[-1][0] + got[-1][1] == spp
            counts = [n for _, n in got]
            assert max(counts) - min(counts) <= 1
    with pytest.raises(ValueError):
        M.sample_range(2, 2, 4)


def test_group_api_host_side():
    """The parts of the ff_group API that need no device: the sample split equals its Python statement, and bad arguments
    come back as status codes (never a crash)."""
    import ctypes as C
    import gpupathtracer_b200 as ff
    for spp in (0, 1, 2, 7, 64, 1000):
        for world in (1, 2, 3, 4, 8):
            for r in range(world):
                a, b = C.c_int(-1), C.c_int(-1)
                assert ff.lib.ff_group_sample_range(r, world, spp, C.byref(a), C.byref(b)) == 0
                assert (a.value, b.value) == M.sample_range(r, world, spp)
    a, b = C.c_int(), C.c_int()
    assert ff.lib.ff_group_sample_range(2, 2, 4, C.byref(a), C.byref(b)) < 0
    h = C.c_void_p()
    assert ff.lib.ff_group_create_rank(None, None, 0, 1, C.byref(h)) < 0 and not h.value
    assert ff.lib.ff_group_create(None, 2, C.byref(h)) < 0 and not h.value
    ids = (C.c_int * 2)(0, 0)
    assert ff.lib.ff_group_create(ids, 2, C.byref(h)) < 0 and b"twice" in ff.lib.ff_last_error()
    assert ff.lib.ff_group_world(None) == 0 and ff.lib.ff_group_rank(None, 0) == -1
    ff.lib.ff_group_destroy(None)


def test_tile_owner_interleave():
    owners = {M.tile_owner(x, y, 4, 32, tiles_x=8) for x in range(0, 256, 32) for y in range(0, 64, 32)}
    assert owners == {0, 1, 2, 3}
    assert M.tile_owner(0, 0, 4, 32, 8) == 0 and M.tile_owner(33, 0, 4, 32, 8) == 1 and M.tile_owner(0, 32, 4, 32, 8) == 0


def test_tile_owned_pixels_partition_the_frame():
    """ff_tile_owned_pixels (the host side of the tile split): ranks' pixel lists are disjoint, cover the frame, agree with
    tile_owner, and are ragged-size safe (frames that are not a multiple of the tile, more ranks than tiles)."""
    import gpupathtracer_b200 as ff
    for (w, h, tile) in ((64, 64, 32), (100, 37, 32), (33, 65, 16), (7, 5, 32), (256, 128, 0)):
        for world in (1, 2, 3, 4, 8):
            seen = np.zeros(w * h, np.int32)
            t = tile or 32
            tiles_x = (w + t - 1) // t
            for r in range(world):
                own = ff.tile_owned_pixels(w, h, r, world, tile)
                seen[own] += 1
                xs, ys = own % w, own // w
                assert all(M.tile_owner(int(x), int(y), world, t, tiles_x) == r for x, y in zip(xs[::17], ys[::17]))
            assert (seen == 1).all(), (w, h, tile, world)
    with pytest.raises(ff.FireflyError):
        ff.tile_owned_pixels(8, 8, 2, 2)


def test_combine_sums_matches_sequential():
    rng = np.random.default_rng(0)
    parts = [rng.uniform(0, 3, (5, 7, 4)).astype(np.float32) for _ in range(4)]
    out = M.combine_sums(parts, 16)
    want = ((((parts[0] + parts[1]) + parts[2]) + parts[3]) * np.float32(1 / 16)).astype(np.float32)
    assert (out[..., :3] == want[..., :3]).all() and (out[..., 3] == 1).all()


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, out_dir):
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist
    from gpupathtracer_b200 import multigpu as MM
    from gpupathtracer_b200 import scenes as S
    from oracle import oracle as O
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    orc = O.Oracle()
    meshes = S.load_meshes()
    desc = S.config2(meshes, 24, 24, 6, 3)
    sc = S.instantiate_on(desc, orc.scene(), meshes)
    spp = desc["spp"]
    first, count = MM.sample_range(rank, world, spp)
    img, _ = sc.render(desc["camera"], count, desc["max_bounces"], first_sample=first, threads=2)
    local_sum = torch.from_numpy((img * np.float32(count)).astype(np.float32))  # the oracle returns means
    MM.reduce_sums(local_sum, dist, 0)
    if rank == 0:
        np.save(os.path.join(out_dir, "reduced.npy"), local_sum.numpy())
    # tile split: every rank keeps only the pixels of its interleaved tiles (all samples), zeros elsewhere; same reduce
    import gpupathtracer_b200 as ff
    full, _ = sc.render(desc["camera"], spp, desc["max_bounces"], threads=2)
    cam = desc["camera"]
    own = ff.tile_owned_pixels(cam["width"], cam["height"], rank, world, 8)
    part = np.zeros_like(full).reshape(-1, full.shape[-1])
    part[own] = full.reshape(-1, full.shape[-1])[own]
    tile_sum = torch.from_numpy(part.reshape(full.shape).copy())
    MM.reduce_sums(tile_sum, dist, 0)
    if rank == 0:
        np.save(os.path.join(out_dir, "tiles.npy"), tile_sum.numpy())
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_sample_split_over_gloo(orc, meshes, tmp_path):
    import torch.multiprocessing as mp
    from gpupathtracer_b200 import scenes as S
    world, port = 2, _free_port()
    mp.spawn(_worker, args=(world, port, str(tmp_path)), nprocs=world, join=True)
    reduced = np.load(tmp_path / "reduced.npy")
    desc = S.config2(meshes, 24, 24, 6, 3)
    sc = S.instantiate_on(desc, orc.scene(), meshes)
    full, _ = sc.render(desc["camera"], desc["spp"], desc["max_bounces"])
    got = reduced / np.float32(desc["spp"])
    # same set of paths; only the fp32 summation order differs (per-rank partial sums, then rank order)
    assert np.abs(got[..., :3] - full[..., :3]).max() <= 1e-5 * max(1.0, float(full[..., :3].max()))
    assert full[..., :3].max() > 0
    # tile split: disjoint pixels, so the reduce reproduces the single-process image bit for bit
    tiles = np.load(tmp_path / "tiles.npy")
    assert (tiles[..., :3].view(np.uint32) == full[..., :3].view(np.uint32)).all()
