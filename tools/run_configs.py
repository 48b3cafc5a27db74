#!/usr/bin/env python3
"""Runs the five BASELINE.json configs on one GPU and prints one JSON line per measurement -- the rows of the results
table in BASELINE.md section 5.  Not the contract bench (that is bench.py); a reproducible table generator.

    python tools/run_configs.py [--configs c1,c2,c3,c4,c5] [--c5-tris 20000000] [--c5-spp 8] [--c3-spp 256] [--no-oracle]

For C1-C3 the primary rays of the frame are also traced by the CPU oracle (the reference's brute-force intersectRays on
all host threads) and compared id-for-id and bit-for-bit in t: `hit_id_mismatches`, and `cpu_mrays_s` is the CPU
baseline measured in the same run.  C4/C5 use the Philox-chosen 65 536-ray subset for that.
C5 at 1024 spp is ~34 G paths; one GPU renders `--c5-spp` of them (throughput does not depend on the sample count)."""
import argparse
import json
import os
import sys
import time

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


def emit(**kw):
    print(json.dumps(kw), flush=True)


def bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


def oracle_check(orc, desc, meshes, sc, rays, what):
    """CPU oracle (brute force, all host threads) on `rays`; returns (mismatches, CPU Mrays/s, cores)."""
    osc = S.instantiate_on(desc, orc.scene(), meshes)
    t0 = time.perf_counter()
    want = osc.intersect(rays, mode=1)
    dt = time.perf_counter() - t0
    got = sc.trace_closest(rays)
    bad = int(((got["geom"] != want["geom"]) | (got["tri"] != want["tri"]) | (bits(got["t"]) != bits(want["t"]))).sum())
    return bad, len(rays) / dt / 1e6, int(orc.threads)


def render_rows(ctx, name, desc, meshes, orc, spp, builders=(0,), nee=(0,)):
    for builder in builders:
        sc, cam = S.instantiate(desc, ctx, meshes, builder=builder)
        info = sc.info()
        base = dict(cfg=name, triangles=info["n_triangles"], nodes=info["n_nodes"], builder=["lbvh", "sah"][builder],
                    build_ms=info["build_ms"], image=[cam.width, cam.height], spp=spp, max_bounces=desc["max_bounces"])
        if orc is not None and builder == 0:
            rays = ctx.camera_rays(cam)
            bad, cpu, cores = oracle_check(orc, desc, meshes, sc, rays, name)
            emit(**base, measure="primary_rays_vs_oracle", rays=len(rays), hit_id_mismatches=bad, cpu_mrays_s=cpu, cpu_cores=cores)
        for n in nee:
            for integ, nm in ((ff.INTEGRATOR_WAVEFRONT, "wavefront"), (ff.INTEGRATOR_MEGAKERNEL, "megakernel")):
                sc.render(cam, min(spp, 4), desc["max_bounces"], host=False, integrator=integ, nee=n)  # warm-up
                ctx.flush_l2()
                _, st = sc.render(cam, spp, desc["max_bounces"], want_stats=True, host=False, integrator=integ, nee=n)
                row = dict(base, measure="render", integrator=nm, nee=n, gpu_ms=st["gpu_ms"], rays=st["rays"], shadow_rays=st["shadow_rays"],
                           mrays_s=(st["rays"] + st["shadow_rays"]) / st["gpu_ms"] / 1e3, msamples_s=st["paths"] / st["gpu_ms"] / 1e3)
                if integ == ff.INTEGRATOR_WAVEFRONT and n == 0:
                    _, c = sc.render(cam, min(spp, 16), desc["max_bounces"], want_stats=True, host=False, integrator=integ | ff.TRACE_COUNT)
                    nn, nt = c["nodes_visited"] / c["rays"], c["tris_tested"] / c["rays"]
                    bpr = 48 + 80 * nn + 48 * nt
                    row.update(trace_ms=st["trace_ms"], nodes_per_ray=nn, tris_per_ray=nt, bytes_per_ray=bpr,
                               frac_of_measured_hbm=bpr * st["rays"] / (st["trace_ms"] * 1e-3) / 1e9 / PEAK if st["trace_ms"] else None)
                emit(**row)
        sc.close()


def soup_rows(ctx, name, n_tris, orc, n_rays, builders):
    soup = ctx.make_soup(n_tris)
    for builder in builders:
        sc = ctx.scene()
        sc.add_triangles(soup)
        sc.commit(builder)
        sc.commit(builder)  # second build: allocator warm
        info = sc.info()
        base = dict(cfg=name, triangles=n_tris, nodes=info["n_nodes"], builder=["lbvh", "sah"][builder], build_ms=info["build_ms"],
                    sah_cost=info["sah_cost"], depth=info["depth"])
        if orc is not None and builder == 0:
            sub = ctx.make_rays_uniform(0, 65536)
            osc = orc.scene()
            osc.add_mesh(soup)
            t0 = time.perf_counter()
            want = osc.intersect(sub, mode=1)
            dt = time.perf_counter() - t0
            got = sc.trace_closest(sub)
            bad = int(((got["geom"] != want["geom"]) | (got["tri"] != want["tri"]) | (bits(got["t"]) != bits(want["t"]))).sum())
            emit(**base, measure="subset_vs_oracle", rays=len(sub), hit_id_mismatches=bad, cpu_mrays_s=len(sub) / dt / 1e6, cpu_cores=int(orc.threads))
        d_rays, d_hits = ctx.alloc(n_rays * 32), ctx.alloc(n_rays * 16)
        ctx.make_rays_uniform(0, n_rays, device_buffer=d_rays, host=False)

        def timed(d_r, n, what):
            for _ in range(2):
                sc.trace_closest_device(d_r, n, d_hits)
            ms = []
            for _ in range(5):
                ctx.flush_l2()
                ms.append(sc.trace_closest_device(d_r, n, d_hits)["gpu_ms"])
            c = sc.trace_closest_device(d_r, n, d_hits, ff.TRACE_BVH8 | ff.TRACE_COUNT)
            nn, nt = c["nodes_visited"] / n, c["tris_tested"] / n
            bpr = 48 + 80 * nn + 48 * nt
            mean = float(np.mean(ms))
            emit(**base, measure=what, rays=n, ms_mean=mean, ms_best=float(np.min(ms)), mrays_s=n / mean / 1e3, nodes_per_ray=nn, tris_per_ray=nt,
                 bytes_per_ray=bpr, alg_gbs=bpr * n / (mean * 1e-3) / 1e9, frac_of_measured_hbm=bpr * n / (mean * 1e-3) / 1e9 / PEAK,
                 frac_of_8tbs=bpr * n / (mean * 1e-3) / 1e9 / 8000.0)

        timed(d_rays, n_rays, "batchA_uniform_incoherent")
        cam = ff.Camera.from_dict(S.config4_camera())
        npr = cam.width * cam.height
        d_prim, d_ph, d_b = ctx.alloc(npr * 32), ctx.alloc(npr * 16), ctx.alloc(npr * 32)
        ctx.camera_rays(cam, device_buffer=d_prim, host=False)
        sc.trace_closest_device(d_prim, npr, d_ph)
        nb = sc.make_rays_bounce(d_prim, d_ph, npr, 1234, d_b)
        timed(d_b, nb, "batchB_diffuse_bounce")
        for b in (d_rays, d_hits, d_prim, d_ph, d_b):
            b.free()
        sc.close()
    return soup


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--configs", default="c1,c2,c3,c4,c5")
    ap.add_argument("--c3-spp", type=int, default=256)
    ap.add_argument("--c5-tris", type=int, default=20000000)
    ap.add_argument("--c5-spp", type=int, default=8)
    ap.add_argument("--no-oracle", action="store_true")
    a = ap.parse_args()
    cfgs = a.configs.split(",")
    ctx = ff.Context(0)
    meshes = S.load_meshes()
    orc = None
    if not a.no_oracle:
        from oracle import oracle as O
        if not O.Oracle.available():
            O.build()
        orc = O.Oracle()
    if "c1" in cfgs:
        render_rows(ctx, "C1", S.config1(), meshes, orc, 1)
    if "c2" in cfgs:
        render_rows(ctx, "C2", S.config2(meshes), meshes, orc, 64, builders=(0, 1), nee=(0, 1))
    if "c3" in cfgs:
        render_rows(ctx, "C3", S.config3(meshes, spp=a.c3_spp), meshes, orc, a.c3_spp, builders=(0, 1), nee=(0, 1))
    if "c4" in cfgs:
        soup_rows(ctx, "C4", 1000000, orc, 3840 * 2160, (0, 1))
    if "c5" in cfgs:
        soup = ctx.make_soup(a.c5_tris)
        desc = S.config5(soup, spp=a.c5_spp)
        t0 = time.perf_counter()
        sc, cam = S.instantiate(desc, ctx)
        commit_wall = time.perf_counter() - t0
        first_build_ms = sc.info()["build_ms"]  # includes growing the builder's scratch arena (one cudaMalloc of ~110 B/triangle)
        sc.commit(0)                             # second build: arena warm, what a rebuild costs
        info = sc.info()
        base = dict(cfg="C5", triangles=info["n_triangles"], nodes=info["n_nodes"], builder="lbvh", build_ms=info["build_ms"],
                    first_build_ms=first_build_ms, commit_wall_s=commit_wall, depth=info["depth"], image=[cam.width, cam.height], spp=a.c5_spp, max_bounces=desc["max_bounces"])
        for n in (0, 1):
            sc.render(cam, 1, desc["max_bounces"], host=False, nee=n)
            _, st = sc.render(cam, a.c5_spp, desc["max_bounces"], want_stats=True, host=False, nee=n)
            row = dict(base, measure="render", integrator="wavefront", nee=n, gpu_ms=st["gpu_ms"], rays=st["rays"], shadow_rays=st["shadow_rays"],
                       mrays_s=(st["rays"] + st["shadow_rays"]) / st["gpu_ms"] / 1e3, msamples_s=st["paths"] / st["gpu_ms"] / 1e3, trace_ms=st["trace_ms"])
            if n == 0:
                _, c = sc.render(cam, 1, desc["max_bounces"], want_stats=True, host=False, integrator=ff.INTEGRATOR_WAVEFRONT | ff.TRACE_COUNT)
                nn, nt = c["nodes_visited"] / c["rays"], c["tris_tested"] / c["rays"]
                bpr = 48 + 80 * nn + 48 * nt
                row.update(nodes_per_ray=nn, tris_per_ray=nt, bytes_per_ray=bpr,
                           frac_of_measured_hbm=bpr * st["rays"] / (st["trace_ms"] * 1e-3) / 1e9 / PEAK if st["trace_ms"] else None)
            emit(**row)
        # properties at full size (no oracle can brute-force 20M x 8K): BVH8 == BVH2 on 1M incoherent rays, == brute force on 2048
        rays = ctx.make_rays_uniform(0, 1 << 20)
        a8, a2 = sc.trace_closest(rays, ff.TRACE_BVH8), sc.trace_closest(rays, ff.TRACE_BVH2)
        bad = int(((a8["geom"] != a2["geom"]) | (a8["tri"] != a2["tri"]) | (bits(a8["t"]) != bits(a2["t"]))).sum())
        br = sc.trace_closest(rays[:2048], ff.TRACE_BRUTE)
        bad2 = int(((a8["geom"][:2048] != br["geom"]) | (a8["tri"][:2048] != br["tri"]) | (bits(a8["t"][:2048]) != bits(br["t"]))).sum())
        emit(**base, measure="properties", bvh8_vs_bvh2_mismatches=bad, rays=len(rays), bvh8_vs_brute_mismatches=bad2, brute_rays=2048)
        sc.close()


if __name__ == "__main__":
    main()
