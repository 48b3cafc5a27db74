"""Probe: with the guard of self_skip_safe compiled out (-DFF_SELF_SKIP_UNGUARDED), for which ray offsets does skipping the
origin triangle change the image against the CPU restatement (which tests every triangle)?"""
import sys
sys.path.insert(0, '/root/repo')
import numpy as np
import gpupathtracer_b200 as ff
from gpupathtracer_b200 import scenes as S
from oracle import oracle as O
orc = O.Oracle(); meshes = S.load_meshes(); ctx = ff.Context(0)
for name, desc in (("C2", S.config2(meshes, 96, 96, 8, 4)), ("C3", S.config3(meshes, 96, 54, 8, 8)), ("shaded", S.shaded(meshes, 64, 64, 4, 5))):
    sc, cam = S.instantiate(desc, ctx, meshes); osc = S.instantiate_on(desc, orc.scene(), meshes)
    for eps in (1e-4, 1e-5, 1e-6, 1e-7, 1e-8, 0.0):
        want, _ = osc.render(desc["camera"], desc["spp"], desc["max_bounces"], seed=5, ray_eps=eps)
        got = sc.render(cam, desc["spp"], desc["max_bounces"], seed=5, ray_eps=eps)
        bad = int((got.view(np.uint32) != want.view(np.uint32)).any(axis=-1).sum())
        print(name, "ray_eps", eps, "pixels that differ:", bad, "of", got.shape[0] * got.shape[1], flush=True)
