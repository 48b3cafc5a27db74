"""Multi-GPU rendering: a thin Python caller of the library's group API (include/firefly.h `ff_group_*`,
csrc/ff_group.cpp).  The reference is single-GPU (a dead `multi_gpu` flag at utilities.h:484-487 is all it has), so this is
new capability; the north star asks for disjoint sample ranges or interleaved tiles and ONE framebuffer reduce per frame
over NVLink -- split, NCCL communicator, reduce and resolve all live behind the C ABI, so a C++ host (firefly_cli
--gpus N, the reference's own main()) gets the same thing without Python.

Because the Philox stream is a pure function of (pixel, global sample index, bounce), the union of the ranks' samples
is the same set of paths whatever the world size; only the order of the final fp32 sums differs.

    Group.local(device_ids)       one process drives several GPUs (one context + host thread per device)
    Group.from_torch(ctx, dist)   one process per GPU under torchrun: torch.distributed is the rendezvous that carries
                                  the 128-byte NCCL id from rank 0 to the others, nothing else
    Group.single(ctx)             a group of one (the one-GPU base of a scaling series: same code path, no reduce)

Reduce paths (`reduce=`): "nccl" = ncclReduce + resolve kernel; "peer" = the fused reduce-scatter + resolve kernel over
NVLink peer memory (sums in rank order: bit-reproducible for a given world size).
The numpy helpers below (sample_range, tile_owner, combine_sums) are the host statements the tests hold the library to.
"""
import ctypes as C

import numpy as np


def sample_range(rank, world, spp):
    """Contiguous, disjoint, exhaustive split of [0, spp): rank k gets [k*spp//G, (k+1)*spp//G)  (== ff_group_sample_range)."""
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
    """Framebuffer reduce through torch.distributed (gloo in the CPU tests of the host logic). local_sum: torch tensor."""
    dist.reduce(local_sum, dst=dst, op=dist.ReduceOp.SUM)
    return local_sum


_SPLITS = {"samples": 0, "strong": 0, "tiles": 1}
_REDUCES = {"nccl": 0, "peer": 1}


class Group:
    """ff_group.  `render` is collective for a from_torch group: every rank calls it with its own committed scene."""

    def __init__(self, handle, contexts, owns_contexts):
        self.h, self.contexts, self._owns = handle, contexts, owns_contexts

    @classmethod
    def single(cls, ctx):
        import gpupathtracer_b200 as ff
        h = C.c_void_p()
        ff._check(ff.lib.ff_group_create_rank(ctx.h, None, 0, 1, C.byref(h)))
        return cls(h, [ctx], False)

    @classmethod
    def from_torch(cls, ctx, dist):
        """One rank per process: rank 0's ncclUniqueId reaches the others through torch.distributed (object broadcast, works
        on nccl and gloo process groups); the communicator itself is created and owned by the library."""
        import gpupathtracer_b200 as ff
        rank, world = dist.get_rank(), dist.get_world_size()
        blob = [None]
        if rank == 0:
            buf = (C.c_ubyte * ff.GROUP_ID_BYTES)()
            ff._check(ff.lib.ff_group_unique_id(buf))
            blob = [bytes(buf)]
        if world > 1:
            dist.broadcast_object_list(blob, src=0)
        h = C.c_void_p()
        idbuf = (C.c_ubyte * ff.GROUP_ID_BYTES).from_buffer_copy(blob[0]) if world > 1 else None
        ff._check(ff.lib.ff_group_create_rank(ctx.h, idbuf, rank, world, C.byref(h)))
        return cls(h, [ctx], False)

    @classmethod
    def local(cls, device_ids):
        """One process, several GPUs.  The contexts belong to the group: build one scene per context (same content)."""
        import gpupathtracer_b200 as ff
        ids = (C.c_int * len(device_ids))(*device_ids)
        h = C.c_void_p()
        ff._check(ff.lib.ff_group_create(ids, len(device_ids), C.byref(h)))
        ctxs = []
        for i in range(len(device_ids)):
            c = ff.Context.__new__(ff.Context)
            c.h, c.device, c._borrowed = C.c_void_p(ff.lib.ff_group_context(h, i)), device_ids[i], True
            ctxs.append(c)
        return cls(h, ctxs, True)

    @property
    def world(self):
        import gpupathtracer_b200 as ff
        return ff.lib.ff_group_world(self.h)

    @property
    def rank(self):
        import gpupathtracer_b200 as ff
        return ff.lib.ff_group_rank(self.h, 0)

    def close(self):
        import gpupathtracer_b200 as ff
        if self.h:
            ff.lib.ff_group_destroy(self.h)
            self.h = C.c_void_p()
            if self._owns:
                for c in self.contexts:
                    c.h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def render(self, scenes, cam, spp, max_bounces, split="samples", reduce="peer", to_host=True, pinned=False, device_out=None, **kw):
        """One frame of `spp` samples per pixel IN TOTAL over the group; returns (image on the root or None, stats dict of the
        first local context + reduce_ms).  scenes: one committed Scene per local context (a single Scene for one context)."""
        import gpupathtracer_b200 as ff
        scenes = list(scenes) if isinstance(scenes, (list, tuple)) else [scenes]
        nl = len(self.contexts)
        assert len(scenes) == nl, "one scene per local context"
        p = ff.render_params(spp, max_bounces, **kw)
        root = self.rank == 0
        out = None
        if root and to_host:
            out = self.contexts[0].pinned_array((cam.height, cam.width, 4)) if pinned else np.empty((cam.height, cam.width, 4), np.float32)
        d_out = None
        if root and device_out is not None:
            d_out = device_out.ptr if isinstance(device_out, ff.DeviceBuffer) else C.c_void_p(int(device_out))
        if root and out is None and d_out is None:  # frame stays on the root's device, in the group's own buffer
            d_out = self._frame_buffer(cam)
        handles = (C.c_void_p * nl)(*[s.h for s in scenes])
        st = (ff.CStats * nl)()
        rms = C.c_double(0)
        ff._check(ff.lib.ff_group_render(self.h, handles, C.byref(cam.c), C.byref(p), _SPLITS[split], _REDUCES[reduce], ff._p(out), d_out,
                                         st, C.byref(rms)))
        stats = st[0].as_dict()
        stats["reduce_ms"] = rms.value
        stats["per_context"] = [st[i].as_dict() for i in range(nl)]
        return out, stats

    def _frame_buffer(self, cam):
        key = (cam.width, cam.height)
        if getattr(self, "_fb_key", None) != key:
            self._fb = self.contexts[0].alloc(cam.width * cam.height * 16)
            self._fb_key = key
        return self._fb.ptr


def render_distributed(scene, cam, spp, max_bounces, dist=None, mode="nccl", split="strong", to_host=True, pinned=False, group=None, **kw):
    """Renders one frame over dist's world through the library's group API; returns (image on rank 0 or None, stats dict).

    split="strong": `spp` is the frame's total, rank k renders sample_range(k, G, spp).
    split="weak":   every rank renders `spp` samples, rank k the global indices [k*spp, (k+1)*spp); total = G*spp.
    split="tiles":  interleaved 32x32 tiles, every rank renders ALL `spp` samples of its tiles -- bit-identical to one GPU.
    Pass `group` to reuse a Group across frames (creating one costs an NCCL communicator)."""
    own = group is None
    if own:
        group = Group.from_torch(scene.ctx, dist) if dist is not None else Group.single(scene.ctx)
    try:
        total = spp * group.world if split == "weak" else spp
        img, st = group.render(scene, cam, total, max_bounces, split="tiles" if split == "tiles" else "samples", reduce=mode, to_host=to_host,
                               pinned=pinned, **kw)
        if img is not None and pinned:
            img = np.array(img)
    finally:
        if own:
            group.close()
    return img, st


def render_distributed_weak(scene, cam, spp, max_bounces, dist, mode="nccl", to_host=True, pinned=False, **kw):
    return render_distributed(scene, cam, spp, max_bounces, dist, mode, "weak", to_host, pinned, **kw)
