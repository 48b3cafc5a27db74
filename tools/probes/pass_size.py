#!/usr/bin/env python3
"""Development probe: wavefront frame time of C2 / C3 as a function of samples per pass (paths in flight)."""
import json, os, sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))
import gpupathtracer_b200 as ff
from gpupathtracer_b200 import scenes as S
ctx = ff.Context(0)
meshes = S.load_meshes()
for name, desc in (("c2", S.config2(meshes)), ("c3", S.config3(meshes, spp=64))):
    sc, cam = S.instantiate(desc, ctx, meshes, builder=1)
    for spb in (4, 8, 16, 32, 64):
        if spb > desc["spp"]:
            continue
        best = 1e30
        for _ in range(4):
            _, st = sc.render(cam, desc["spp"], desc["max_bounces"], want_stats=True, host=False, samples_per_pass=spb)
            best = min(best, st["gpu_ms"])
        print(json.dumps(dict(case=name, samples_per_pass=spb, gpu_ms=best, trace_ms=st["trace_ms"], mrays_s=st["rays"] / best / 1e3)), flush=True)
    sc.close()
