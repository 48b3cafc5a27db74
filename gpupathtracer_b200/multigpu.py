"""Multi-GPU rendering: one process per GPU (torchrun), scene + BVH replicated, samples split across ranks, one
framebuffer reduce per frame.  The reference is single-GPU (a dead `multi_gpu` flag at
utilities.h:484-487 is all it has), so this is new capability; the north star asks for disjoint sample ranges or
interleaved tiles and ONE reduce per frame over NVLink.

Because the Philox stream is a pure function of (pixel, global sample index, bounce), the union of the ranks' samples
is the same set of paths whatever the world size; only the order of the final fp32 sums differs.

Two reduce paths:
  * "nccl":  torch.distributed.reduce(SUM) of the RGBA fp32 accumulator to rank 0, then ff_resolve   (baseline)
  * "peer":  every rank publishes its accumulator through a CUDA IPC handle; rank 0 maps them and ONE kernel reads
             the peers over NVLink, adds them in rank order and applies 1/spp (ff_reduce_resolve_peers) -- the reduce
             and the resolve are fused, and the result is deterministic for a given world size.
torch.distributed is plumbing only (rendezvous, barrier, the baseline reduce, handle exchange).
"""
import numpy as np


def sample_range(rank, world, spp):
    """Contiguous, disjoint, exhaustive split of [0, spp): rank k gets [k*spp//G, (k+1)*spp//G)."""
    if world <= 0 or not (0 <= rank < world) or spp < 0:
        raise ValueError("bad rank/world/spp")
    a = rank * spp // world
    b = (rank + 1) * spp // world
    return a, b - a


def tile_owner(x, y, world, tile=32, tiles_x=None):
    """Interleaved screen tiles: tile t (row-major) belongs to rank t mod world."""
    if tiles_x is None:
        raise ValueError("tiles_x required")
    return ((y // tile) * tiles_x + (x // tile)) % world


def combine_sums(partials, total_spp):
    """Host statement of what the reduce + resolve computes: sums added in rank order, times fp32(1/total_spp)."""
    acc = np.array(partials[0], np.float32, copy=True)
    for p in partials[1:]:
        acc = (acc + np.asarray(p, np.float32)).astype(np.float32)
    out = (acc * np.float32(1.0 / np.float32(total_spp))).astype(np.float32)
    out[..., 3] = 1.0
    return out


def reduce_sums(local_sum, dist, dst=0):
    """Framebuffer reduce through torch.distributed (NCCL on GPUs, gloo in the CPU tests). local_sum: torch tensor."""
    dist.reduce(local_sum, dst=dst, op=dist.ReduceOp.SUM)
    return local_sum


def render_distributed(scene, cam, spp, max_bounces, dist=None, mode="nccl", split="strong", to_host=True, pinned=False, **kw):
    """Renders one frame over dist's world; returns (image on rank 0 or None, this rank's ff stats dict).

    split="strong": `spp` is the frame's total, rank k renders sample_range(k, G, spp).
    split="weak":   every rank renders `spp` samples, rank k the global indices [k*spp, (k+1)*spp); total = G*spp.
    split="tiles":  every rank renders ALL `spp` samples of the pixels of its interleaved 32x32 tiles (tile_owner); the other
                    pixels of its accumulator stay zero, so the same SUM reduce assembles the frame -- bit-identical to a
                    single-GPU render for every world size (x + 0 = x), the split for low-spp / interactive frames.
    `scene` must already be committed on this rank's GPU (scene + BVH are replicated).  With to_host=False the resolved
    frame stays in device memory on rank 0 (returned image is None everywhere)."""
    import ctypes as C
    import torch
    import gpupathtracer_b200 as ff
    rank = dist.get_rank() if dist is not None else 0
    world = dist.get_world_size() if dist is not None else 1
    if split == "tiles":
        first, count, total = 0, spp, spp
        if world > 1:
            kw = dict(kw, tile_rank=rank, tile_world=world, tile_size=kw.get("tile_size", 32))
    elif split == "weak":
        first, count, total = rank * spp, spp, spp * world
    else:
        (first, count), total = sample_range(rank, world, spp), spp
    w, h = cam.width, cam.height
    dev = torch.device("cuda", scene.ctx.device)
    peer = world > 1 and mode == "peer"
    if peer:  # a cudaMalloc base pointer of our own: an IPC handle names a whole allocation, not a sub-block of torch's pool
        own = scene.ctx.alloc(h * w * 16)
        own.zero()
        accum_ptr = own.ptr.value
    else:
        accum = torch.zeros((h, w, 4), dtype=torch.float32, device=dev)
        torch.cuda.synchronize(dev)
        accum_ptr = accum.data_ptr()
    stats = dict(gpu_ms=0.0, rays=0, paths=0, kernel_launches=0, trace_ms=0.0)
    if count > 0:
        stats = scene.render_accumulate(cam, accum_ptr, count, max_bounces, first_sample=first, **kw)
    d_out = None
    if not to_host:
        d_out = torch.empty((h, w, 4), dtype=torch.float32, device=dev)
    img = None
    if world == 1 or mode == "nccl":
        if world > 1:
            reduce_sums(accum, dist, 0)
            torch.cuda.synchronize(dev)
        if rank == 0:
            img = scene.ctx.resolve(accum_ptr, w, h, total, host=to_host, device_out=d_out.data_ptr() if d_out is not None else None, pinned=pinned)
        return img, stats
    if mode == "peer":
        handle = np.zeros(64, np.uint8)
        ff._check(ff.lib.ff_ipc_get_handle(scene.ctx.h, C.c_void_p(accum_ptr), handle.ctypes.data_as(C.c_void_p)))
        mine = torch.from_numpy(handle.copy()).to(dev)
        gathered = [torch.zeros(64, dtype=torch.uint8, device=dev) for _ in range(world)]
        dist.all_gather(gathered, mine)
        if rank == 0:
            ptrs, opened = [], []
            for r in range(world):
                if r == 0:
                    ptrs.append(accum_ptr)
                    continue
                hb = np.ascontiguousarray(gathered[r].cpu().numpy())
                p = C.c_void_p()
                ff._check(ff.lib.ff_ipc_open_handle(scene.ctx.h, hb.ctypes.data_as(C.c_void_p), C.byref(p)))
                ptrs.append(p.value)
                opened.append(p)
            img = scene.ctx.reduce_resolve_peers(ptrs, w, h, total, host=to_host, device_out=d_out.data_ptr() if d_out is not None else None, pinned=pinned)
            for p in opened:
                ff.lib.ff_ipc_close_handle(scene.ctx.h, p)
        dist.barrier()  # peers keep their accumulators alive until rank 0 has read them
        own.free()
        return img, stats
    raise ValueError("mode must be 'nccl' or 'peer'")


def render_distributed_weak(scene, cam, spp, max_bounces, dist, mode="nccl", to_host=True, pinned=False, **kw):
    return render_distributed(scene, cam, spp, max_bounces, dist, mode, "weak", to_host, pinned, **kw)
