#!/usr/bin/env python3
"""Multi-GPU parity check, one rank per GPU:

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P tools/multigpu_check.py

Every rank commits the same scene, renders its share of the samples, and the framebuffers are combined by one reduce per
frame (NCCL reduce, and the fused peer-memory reduce+resolve kernel).  Rank 0 compares both against a single-GPU render
of the whole frame: the set of paths is identical (Philox is keyed by the global sample index), only the association of
the per-rank partial sums differs, so the bar is 1e-5 absolute.  The peer path must also equal the host statement of
"sum in rank order, times 1/spp" bit for bit.  Prints one JSON line on rank 0 and exits non-zero on a mismatch."""
import json
import os
import sys

import numpy as np

ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
sys.path.insert(0, ROOT)


def main():
    import torch
    import torch.distributed as dist
    rank, local, world = int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    import gpupathtracer_b200 as ff
    from gpupathtracer_b200 import multigpu as M
    from gpupathtracer_b200 import scenes as S
    ctx = ff.Context(local)
    meshes = S.load_meshes()
    w, h, spp, bounces = 160, 128, 8 * world, 3
    desc = S.config2(meshes, w, h, spp, bounces)
    sc, cam = S.instantiate(desc, ctx, meshes)
    out = dict(world=world, image=[w, h], spp=spp)
    ok = True
    full = sc.render(cam, spp, bounces) if rank == 0 else None
    for mode in ("nccl", "peer"):
        for split in ("strong", "weak"):
            n = spp if split == "strong" else spp // world
            img, st = M.render_distributed(sc, cam, n, bounces, dist, mode=mode, split=split)
            if rank == 0:
                err = float(np.abs(img - full).max())
                out["%s_%s_max_abs_err" % (mode, split)] = err
                ok = ok and err <= 1e-5 and bool((img[..., 3] == 1.0).all())
    # interleaved-tile split: disjoint pixels per rank, so both reduce paths must reproduce the single-GPU frame bit for bit
    for mode in ("nccl", "peer"):
        img, st = M.render_distributed(sc, cam, spp, bounces, dist, mode=mode, split="tiles")
        if rank == 0:
            same = bool((img[..., :3].view(np.uint32) == full[..., :3].view(np.uint32)).all())
            out["%s_tiles_bit_identical" % mode] = same
            ok = ok and same
    # the peer kernel against its host statement: partial sums gathered through NCCL, added in rank order
    first, count = M.sample_range(rank, world, spp)
    acc = torch.zeros((h, w, 4), dtype=torch.float32, device="cuda")
    sc.render_accumulate(cam, acc.data_ptr(), count, bounces, first_sample=first)
    parts = [torch.zeros_like(acc) for _ in range(world)]
    dist.all_gather(parts, acc)
    img, _ = M.render_distributed(sc, cam, spp, bounces, dist, mode="peer", split="strong")
    if rank == 0:
        want = M.combine_sums([p.cpu().numpy() for p in parts], spp)
        same = bool((want.view(np.uint32) == img.view(np.uint32)).all())
        out["peer_bit_identical_to_rank_order_sum"] = same
        ok = ok and same
        out["ok"] = ok
        print(json.dumps(out), flush=True)
    flag = torch.tensor([1 if ok else 0], device="cuda")
    dist.broadcast(flag, 0)
    dist.barrier()
    dist.destroy_process_group()
    return 0 if int(flag.item()) == 1 else 1


if __name__ == "__main__":
    sys.exit(main())
