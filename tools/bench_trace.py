#!/usr/bin/env python3
"""Kernel-level throughput probe used while optimising (not the contract bench): closest-hit Mrays/s with the ray batch
resident in HBM, for config 4 (1M-triangle soup; batch A uniform-incoherent, batch B diffuse-bounce) and for the C2/C3
frames, with nodes/ray, triangles/ray and algorithmic bytes.  Prints one JSON object per case.

    python tools/bench_trace.py [--tris 1000000] [--rays 8294400] [--reps 5] [--cases soupA,soupB,c2,c3]
"""
import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
sys.path.insert(0, ROOT)
import gpupathtracer_b200 as ff  # noqa: E402
from gpupathtracer_b200 import scenes as S  # noqa: E402

PEAK = 6559.0
try:
    PEAK = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
except Exception:
    pass


def time_trace(ctx, sc, d_rays, n, d_hits, mode, reps):
    best, tot = 1e30, 0.0
    for _ in range(2):
        sc.trace_closest_device(d_rays, n, d_hits, mode)
    for _ in range(reps):
        ctx.flush_l2()
        st = sc.trace_closest_device(d_rays, n, d_hits, mode)
        best = min(best, st["gpu_ms"])
        tot += st["gpu_ms"]
    stc = sc.trace_closest_device(d_rays, n, d_hits, mode | ff.TRACE_COUNT)
    nn, nt = stc["nodes_visited"] / n, stc["tris_tested"] / n
    bpr = 48 + 80 * nn + 48 * nt
    mean = tot / reps
    return dict(ms_best=best, ms_mean=mean, mrays_s=n / mean / 1e3, nodes_per_ray=nn, tris_per_ray=nt, bytes_per_ray=bpr,
                alg_gbs=bpr * n / (mean * 1e-3) / 1e9, frac_of_measured_hbm=bpr * n / (mean * 1e-3) / 1e9 / PEAK)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--tris", type=int, default=1000000)
    ap.add_argument("--rays", type=int, default=3840 * 2160)
    ap.add_argument("--reps", type=int, default=5)
    ap.add_argument("--cases", default="soupA,soupB,c2,c3")
    ap.add_argument("--modes", default="bvh8")
    ap.add_argument("--builder", type=int, default=0)
    ap.add_argument("--tune", default="")  # postpone_div,refill_lanes,blocks_per_sm
    a = ap.parse_args()
    cases = a.cases.split(",")
    modes = {"bvh8": ff.TRACE_BVH8, "bvh8s": ff.TRACE_BVH8 | ff.TRACE_STATIC, "bvh2": ff.TRACE_BVH2}
    ctx = ff.Context(0)
    if a.tune:
        ff.lib.ff_debug_set_tuning(*[int(x) for x in a.tune.split(",")])
    if "soupA" in cases or "soupB" in cases:
        soup = ctx.make_soup(a.tris)
        sc = ctx.scene()
        sc.add_triangles(soup)
        sc.commit(a.builder)
        sc.commit(a.builder)
        info = sc.info()
        print(json.dumps(dict(case="soup_build", tris=a.tris, build_ms=info["build_ms"], nodes=info["n_nodes"], depth=info["depth"],
                              sah=info["sah_cost"])), flush=True)
        n = a.rays
        d_rays, d_hits = ctx.alloc(n * 32), ctx.alloc(n * 16)
        if "soupA" in cases:
            ctx.make_rays_uniform(0, n, device_buffer=d_rays, host=False)
            for m in a.modes.split(","):
                print(json.dumps(dict(case="soupA_uniform_incoherent", mode=m, tris=a.tris, rays=n, **time_trace(ctx, sc, d_rays, n, d_hits, modes[m], a.reps))), flush=True)
        if "soupA_sorted" in cases:
            # experiment: the same batch with the rays regrouped on the host by (origin cell, direction octant); tells whether
            # a device-side coherence sort in front of the extend stage would pay for itself
            host = ctx.make_rays_uniform(0, n)
            for bits in (3, 4, 5, 6):
                q = np.minimum((host[:, :3] * (1 << bits)).astype(np.uint32), (1 << bits) - 1)
                key = np.zeros(n, np.uint64)
                for b in range(bits):
                    for ax in range(3):
                        key |= ((q[:, ax] >> b) & 1).astype(np.uint64) << np.uint64(3 * b + ax)
                octant = ((host[:, 3] < 0).astype(np.uint64) << np.uint64(2)) | ((host[:, 4] < 0).astype(np.uint64) << np.uint64(1)) | (host[:, 5] < 0).astype(np.uint64)
                for name, k in (("cell", key), ("cell+octant", (key << np.uint64(3)) | octant), ("octant+cell", (octant << np.uint64(3 * bits)) | key)):
                    order = np.argsort(k, kind="stable")
                    hs = host[order]
                    packed = np.zeros((2 * n, 4), np.float32)
                    packed[:n, :3] = hs[:, :3]
                    packed[n:, :3] = hs[:, 3:]
                    d_rays.upload(packed)
                    print(json.dumps(dict(case="soupA_sorted", key=name, bits=bits, **time_trace(ctx, sc, d_rays, n, d_hits, modes["bvh8"], a.reps))), flush=True)
        if "soupB" in cases:
            cam = ff.Camera.from_dict(S.config4_camera())
            npr = cam.width * cam.height
            d_prim = ctx.alloc(npr * 32)
            d_ph = ctx.alloc(npr * 16)
            ctx.camera_rays(cam, device_buffer=d_prim, host=False)
            sc.trace_closest_device(d_prim, npr, d_ph)
            d_b = ctx.alloc(npr * 32)
            nb = sc.make_rays_bounce(d_prim, d_ph, npr, 1234, d_b)
            for m in a.modes.split(","):
                print(json.dumps(dict(case="soupA_primary_coherent", mode=m, tris=a.tris, rays=npr, **time_trace(ctx, sc, d_prim, npr, d_ph, modes[m], a.reps))), flush=True)
                print(json.dumps(dict(case="soupB_diffuse_bounce", mode=m, tris=a.tris, rays=nb, **time_trace(ctx, sc, d_b, nb, d_hits, modes[m], a.reps))), flush=True)
        sc.close()
    meshes = S.load_meshes()
    for name, desc in (("c2", S.config2(meshes)), ("c3", S.config3(meshes, spp=16))):
        if name not in cases:
            continue
        sc, cam = S.instantiate(desc, ctx, meshes, builder=a.builder)
        for integ, nm in ((ff.INTEGRATOR_WAVEFRONT, "wavefront"), (ff.INTEGRATOR_MEGAKERNEL, "megakernel")):
            for _ in range(2):
                sc.render(cam, 8, desc["max_bounces"], host=False, integrator=integ)
            _, st = sc.render(cam, desc["spp"], desc["max_bounces"], want_stats=True, host=False, integrator=integ)
            out = dict(case=name, integrator=nm, spp=desc["spp"], gpu_ms=st["gpu_ms"], trace_ms=st["trace_ms"], rays=st["rays"],
                       mrays_s=st["rays"] / st["gpu_ms"] / 1e3, msamples_s=st["paths"] / st["gpu_ms"] / 1e3)
            if integ == ff.INTEGRATOR_WAVEFRONT:
                _, sc_ = sc.render(cam, desc["spp"], desc["max_bounces"], want_stats=True, host=False, integrator=integ | ff.TRACE_COUNT)
                out.update(nodes_per_ray=sc_["nodes_visited"] / sc_["rays"], tris_per_ray=sc_["tris_tested"] / sc_["rays"])
            print(json.dumps(out), flush=True)
        sc.close()


if __name__ == "__main__":
    main()
