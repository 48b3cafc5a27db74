#!/usr/bin/env python3
"""Development probe (not a test, not the bench): exercises every GPU path once and prints mismatches / timings."""
import os
import sys
import time

import numpy as np

ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
sys.path.insert(0, ROOT)
import gpupathtracer_b200 as ff  # noqa: E402
from gpupathtracer_b200 import scenes as S  # noqa: E402
from oracle import oracle as O  # noqa: E402


def cmp_hits(name, a, b):
    bad_id = int(((a["geom"] != b["geom"]) | (a["tri"] != b["tri"])).sum())
    both = (a["geom"] >= 0) & (b["geom"] >= 0)
    bad_t = int((a["t"][both] != b["t"][both]).sum())
    print("  %-28s rays %8d hits %8d id-mismatch %d t-bit-mismatch %d" % (name, len(a["geom"]), int((a["geom"] >= 0).sum()), bad_id, bad_t))
    return bad_id, bad_t


def main():
    big = "--big" in sys.argv
    O.build()
    orc = O.Oracle()
    meshes = S.load_meshes()
    ctx = ff.Context(0)
    for desc in (S.config1(), S.config2(meshes, 64, 64, 2, 4), S.config3(meshes, 96, 54, 2, 4), S.reference_main(96, 96)):
        print(desc["name"])
        sc, cam = S.instantiate(desc, ctx, meshes)
        print("  info", sc.info())
        osc = S.instantiate_on(desc, orc.scene(), meshes)
        rays = ctx.camera_rays(cam)
        orays, _ = orc.camera_rays(desc["camera"])
        print("  camera rays bit-equal:", bool((rays.view(np.uint32) == orays.view(np.uint32)).all()))
        ref = osc.intersect(orays, mode=1)
        for mode, nm in ((ff.TRACE_BRUTE, "brute"), (ff.TRACE_BVH2, "bvh2"), (ff.TRACE_BVH8, "bvh8")):
            cmp_hits(nm + " vs oracle", sc.trace_closest(orays, mode), ref)
        if desc["name"] != "main":
            img, st = sc.render(cam, desc["spp"], desc["max_bounces"], want_stats=True)
            oimg, nrays = osc.render(desc["camera"], desc["spp"], desc["max_bounces"])
            print("  wavefront vs oracle: max abs diff %g, bit-equal %s, rays %d vs %d" % (
                float(np.abs(img - oimg).max()), bool((img.view(np.uint32) == oimg.view(np.uint32)).all()), st["rays"], nrays))
            img2, st2 = sc.render(cam, desc["spp"], desc["max_bounces"], want_stats=True, integrator=ff.INTEGRATOR_MEGAKERNEL)
            print("  megakernel vs wavefront bit-equal %s, rays %d" % (bool((img.view(np.uint32) == img2.view(np.uint32)).all()), st2["rays"]))
        u8 = sc.render_u8(cam)
        o8 = osc.render_as_coded(desc["camera"], mode=1)
        print("  as-coded vs oracle(mode 1): differing bytes", int((u8 != o8).sum()))

    for n in ([100000, 1000000] if big else [100000]):
        print("soup", n)
        soup = ctx.make_soup(n)
        osoup = orc.make_soup(n)
        print("  soup bit-equal:", bool((soup.view(np.uint32) == osoup.view(np.uint32)).all()))
        sc = ctx.scene()
        sc.add_triangles(soup)
        t0 = time.time()
        ms = sc.commit()
        print("  commit: device build %.3f ms, wall %.1f ms" % (ms, 1e3 * (time.time() - t0)), sc.info())
        nr = 1 << 20
        rays = ctx.make_rays_uniform(0, nr)
        orays = orc.make_rays_uniform(0, 4096)
        print("  rays bit-equal:", bool((rays[:4096].view(np.uint32) == orays.view(np.uint32)).all()))
        osc = orc.scene()
        osc.add_mesh(soup)
        sub = rays[:2048]
        ref = osc.intersect(sub, mode=1)
        cmp_hits("bvh8 vs oracle (subset)", sc.trace_closest(sub, ff.TRACE_BVH8), ref)
        cmp_hits("brute vs oracle (subset)", sc.trace_closest(sub, ff.TRACE_BRUTE), ref)
        nb = 1 << 16 if n > 200000 else 1 << 18
        cmp_hits("bvh8 vs gpu brute", sc.trace_closest(rays[:nb], ff.TRACE_BVH8), sc.trace_closest(rays[:nb], ff.TRACE_BRUTE))
        cmp_hits("bvh2 vs bvh8", sc.trace_closest(rays, ff.TRACE_BVH2), sc.trace_closest(rays, ff.TRACE_BVH8))
        # device-resident timing
        d_rays = ctx.alloc(nr * 32)
        d_hits = ctx.alloc(nr * 16)
        ctx.make_rays_uniform(0, nr, device_buffer=d_rays, host=False)
        for mode, nm in ((ff.TRACE_BVH8, "bvh8"), (ff.TRACE_BVH2, "bvh2")):
            for _ in range(2):
                st = sc.trace_closest_device(d_rays, nr, d_hits, mode)
            stc = sc.trace_closest_device(d_rays, nr, d_hits, mode | ff.TRACE_COUNT)
            print("  %s: %.3f ms for %d rays = %.1f Mrays/s; nodes/ray %.2f tris/ray %.2f" % (
                nm, st["gpu_ms"], nr, nr / st["gpu_ms"] / 1e3, stc["nodes_visited"] / nr, stc["tris_tested"] / nr))
    print("PROBE DONE")


if __name__ == "__main__":
    main()
