#!/usr/bin/env python3
"""Probe: ff_trace_closest on page-locked host buffers (bench.py's `e2e`) for the streaming path at several chunk sizes and
for the chunked pipeline (FF_TRACE_STREAM=0).  The knobs are read once per process, so every configuration is a subprocess.
    python tools/probes/e2e_stream.py            # the sweep
    python tools/probes/e2e_stream.py --one      # one configuration (environment as given), one JSON line"""
import json
import os
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)


def one():
    import gpupathtracer_b200 as ff
    ctx = ff.Context(0)
    n_tris, n = 1000000, 3840 * 2160
    sc = ctx.scene()
    sc.add_triangles(ctx.make_soup(n_tris))
    sc.commit()
    d_rays, d_hits = ctx.alloc(n * 32), ctx.alloc(n * 16)
    ctx.make_rays_uniform(0, n, device_buffer=d_rays, host=False)
    dev_ms = []
    for i in range(6):
        ctx.flush_l2()
        st = sc.trace_closest_device(d_rays, n, d_hits)
        if i >= 2:
            dev_ms.append(st["gpu_ms"])
    want = sc.hits_unpack(d_hits, n)
    h_rays = ctx.pinned_array((n, 6), np.float32, key="rays")
    h_rays[:] = ctx.make_rays_uniform(0, n)
    h_geom, h_tri, h_t = (ctx.pinned_array((n,), np.int32, key="geom"), ctx.pinned_array((n,), np.int32, key="tri"),
                          ctx.pinned_array((n,), np.float32, key="t"))
    ms, tr, h2d, d2h = [], [], [], []
    for i in range(8):
        h_geom[:] = -7
        ctx.flush_l2()
        ctx.synchronize()
        ctx.timer_begin()
        st = sc.trace_closest_into(h_rays, h_geom, h_tri, h_t)
        t = ctx.timer_end()
        if i >= 2:
            ms.append(t); tr.append(st["trace_ms"]); h2d.append(st["h2d_ms"]); d2h.append(st["d2h_ms"])
    bad = int((want["geom"] != h_geom).sum() + (want["tri"] != h_tri).sum() + (want["t"].view(np.uint32) != h_t.view(np.uint32)).sum())
    # ragged sizes through the same path (one chunk, chunk boundary +- 1, tiny)
    ragged = 0
    for m in (1, 31, 4097, (1 << 18) - 1, (1 << 18) + 1, 777777):
        g = sc.trace_closest(np.array(h_rays[:m]))
        ragged += int((g["geom"] != want["geom"][:m]).sum() + (g["tri"] != want["tri"][:m]).sum() + (g["t"].view(np.uint32) != want["t"][:m].view(np.uint32)).sum())
    # pageable buffers
    g = sc.trace_closest(np.array(h_rays[:1500000]))
    ragged += int((g["geom"] != want["geom"][:1500000]).sum())
    print(json.dumps(dict(stream=os.environ.get("FF_TRACE_STREAM", "1"), shift=os.environ.get("FF_STREAM_CHUNK_SHIFT", "default"),
                          chunk=os.environ.get("FF_TRACE_CHUNK", "default"), preload=os.environ.get("FF_STREAM_PRELOAD", "0"), done_shift=os.environ.get("FF_STREAM_DONE_SHIFT", "same"), device_ms=float(np.mean(dev_ms)), e2e_ms=float(np.mean(ms)),
                          e2e_ms_min=float(np.min(ms)), mrays_s=n / float(np.mean(ms)) / 1e3, trace_ms=float(np.mean(tr)), h2d_ms=float(np.mean(h2d)),
                          d2h_ms=float(np.mean(d2h)), passes=st["passes"], mismatches=bad, ragged_mismatches=ragged)), flush=True)


def main():
    if "--one" in sys.argv:
        return one()
    configs = [dict(FF_TRACE_STREAM="0"), dict(FF_TRACE_STREAM="1"), dict(FF_TRACE_STREAM="1"), dict(FF_TRACE_STREAM="1", FF_STREAM_PRELOAD="1")]
    for cfg in configs:
        env = dict(os.environ)
        env.update(cfg)
        r = subprocess.run([sys.executable, os.path.abspath(__file__), "--one"], env=env, capture_output=True, text=True, timeout=300)
        sys.stdout.write(r.stdout if r.returncode == 0 else "FAILED %s rc=%d\n%s\n" % (cfg, r.returncode, r.stderr[-1500:]))
        sys.stdout.flush()


if __name__ == "__main__":
    main()
