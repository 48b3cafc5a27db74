#!/usr/bin/env python3
"""BASELINE.json config 5: synthetic 20 M-triangle soup + one emissive quad, 7680x4320, `--spp` samples per pixel split
over the ranks as disjoint sample ranges (strong scaling: the frame is fixed), ONE framebuffer reduce per frame.

    python tools/bench_c5.py [--spp 1024] [--tris 20000000] [--reduce nccl|peer] [--nee 0|1] [--out frame.png]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P tools/bench_c5.py ...

Every rank builds its own copy of the scene and BVH (replicated).  Timing: CUDA events on the library's
stream around the whole frame (render + reduce + resolve), barrier + synchronize on both sides, max over ranks.  Rank 0
prints one JSON line: samples/s over all GPUs, per GPU, build time, reduce share."""
import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--spp", type=int, default=1024)
    ap.add_argument("--tris", type=int, default=20000000)
    ap.add_argument("--width", type=int, default=7680)
    ap.add_argument("--height", type=int, default=4320)
    ap.add_argument("--bounces", type=int, default=4)
    ap.add_argument("--reduce", default="nccl", choices=["nccl", "peer"])
    ap.add_argument("--nee", type=int, default=0)
    ap.add_argument("--builder", type=int, default=0)
    ap.add_argument("--out", default="")
    a = ap.parse_args()
    rank, local, world = int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    real_stdout = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)  # NCCL prints its banner on fd 1
    dist = None
    if world > 1:
        import torch
        import torch.distributed as td
        torch.cuda.set_device(local)
        td.init_process_group("nccl", device_id=torch.device("cuda", local))
        dist = td
    import gpupathtracer_b200 as ff
    from gpupathtracer_b200 import multigpu as M
    from gpupathtracer_b200 import scenes as S
    ctx = ff.Context(local)
    soup = ctx.make_soup(a.tris)
    desc = S.config5(soup, a.width, a.height, a.spp, a.bounces)
    sc, cam = S.instantiate(desc, ctx, builder=a.builder)
    sc.commit(a.builder)  # second build: scratch arena warm, the steady-state figure
    info = sc.info()
    del soup

    def barrier():
        ctx.synchronize()
        if dist is not None:
            import torch
            torch.cuda.synchronize()
            dist.barrier()

    # one group for both frames (creating one costs an NCCL communicator); warm-up frame at 1 spp per rank (allocations, NCCL channels)
    group = M.Group.from_torch(ctx, dist) if dist is not None else M.Group.single(ctx)
    M.render_distributed(sc, cam, world, a.bounces, dist, mode=a.reduce, split="strong", to_host=False, nee=a.nee, group=group)
    barrier()
    ctx.timer_begin()
    img, st = M.render_distributed(sc, cam, a.spp, a.bounces, dist, mode=a.reduce, split="strong", to_host=bool(a.out), nee=a.nee, group=group)
    barrier()
    ms = ctx.timer_end()
    vals = [ms, st["gpu_ms"], float(st["rays"] + st.get("shadow_rays", 0)), float(st["paths"])]
    if dist is not None:
        import torch
        t = torch.tensor(vals, dtype=torch.float64, device="cuda")
        mx = t.clone()
        dist.all_reduce(mx, op=dist.ReduceOp.MAX)
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        ms, render_ms, rays, paths = float(mx[0]), float(mx[1]), float(t[2]), float(t[3])
    else:
        ms, render_ms, rays, paths = vals
    if rank == 0:
        npix = a.width * a.height
        line = dict(cfg="C5", n_gpus=world, triangles=info["n_triangles"], bvh_nodes=info["n_nodes"], builder=["lbvh", "sah"][a.builder],
                    build_ms=info["build_ms"], image=[a.width, a.height], total_spp=a.spp, max_bounces=a.bounces, nee=a.nee, reduce=a.reduce,
                    frame_ms=ms, render_ms_max_rank=render_ms, reduce_and_resolve_ms=ms - render_ms, framebuffer_bytes_per_rank=npix * 16,
                    msamples_s=paths / ms / 1e3, msamples_s_per_gpu=paths / ms / 1e3 / world, mrays_s=rays / ms / 1e3, scaling="strong")
        real_stdout.write(json.dumps(line) + "\n")
        real_stdout.flush()
        if a.out and img is not None:
            ff.write_image(a.out, img)
    group.close()
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
