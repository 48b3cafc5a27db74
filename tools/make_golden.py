#!/usr/bin/env python3
"""Generates tests/golden/*.npz from the REFERENCE ITSELF (oracle/_ref = the reference's own routine text compiled
from /root/reference by oracle/Makefile).  Run in the build container, where /root/reference is mounted:

    make -C oracle && python tools/make_golden.py

The reference ships no tests or vectors of its own, so these files are what pins the oracle
restatement (tests/test_oracle_golden.py) and, through it, the CUDA path.  Nothing here is read at run time on
the GPU box except the produced .npz files.

  meshes.npz         positions of the six sceneResources/*.obj as float32 [n,9], loaded by the reference's own
                     tinyobjloader path (ref_load_obj = guarded LoadMesh, utilities.h:781-840)
  kat_triangle.npz   known-answer rays against single triangles through the reference's intersectTriangle
  ref_identity.npz   reference intersectRays (object-space path, kernel.cu:127-176) on identity-transform scenes:
                     bit-exact contract for hit ids AND t
  ref_c1.npz         config 1 (cube + translated triangle), 256x256 camera rays, reference results
  ref_main.npz       the main() scene (rocketman rot (0,90,180) + 4 planes x5): ids, t, and the as-coded 8-bit frame
  ref_scaled.npz     meshes under scales below / above 1 (scenes.scaled_meshes): ids, t and the as-coded 8-bit frame, whose
                     colour is the un-normalised inverse-transpose normal and wraps above 255 (kernel.cu:117,214)
  ref_camera.npz     camera basis, inverse matrices and rays for three cameras
  ref_matrices.npz   model / inverse matrices of Geometry for a set of TRS triples
"""
import os
import sys

import numpy as np

ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
sys.path.insert(0, ROOT)
from oracle import oracle as O  # noqa: E402
from gpupathtracer_b200 import scenes as S  # noqa: E402  (pure-python scene descriptions; needs the built library only for instantiate)

OUT = os.path.join(ROOT, "tests", "golden")
REF_OBJ = "/root/reference/sceneResources"


def main():
    os.makedirs(OUT, exist_ok=True)
    O.build()
    ref = O.Ref()
    rng = np.random.default_rng(20261104)
    only = set(sys.argv[1:])  # e.g. `python tools/make_golden.py ref_scaled.npz`: write just that file (everything is still computed)

    def save(fname, **arrays):
        if only and fname not in only:
            print("not rewritten:", fname)
            return
        np.savez_compressed(os.path.join(OUT, fname), **arrays)

    meshes = {k: ref.load_obj(os.path.join(REF_OBJ, k + ".obj")) for k in S.MESH_NAMES}
    save("meshes.npz", **meshes)
    print({k: v.shape for k, v in meshes.items()})

    # ---- single-triangle known answers
    tri0 = meshes["triangle"][0]
    cases = []

    def add(v9, o, d):
        hit, t, p, n = ref.intersect_triangle(v9, o, d)
        cases.append((np.array(v9, np.float32), np.array(o, np.float32), np.array(d, np.float32), hit, t, p, n))

    add(tri0, (0.25, 0.25, 1), (0, 0, -1))       # front face, t = 1
    add(tri0, (0.25, 0.25, -1), (0, 0, 1))       # back face: culled
    add(tri0, (0.5, 0.5, 1), (0, 0, -1))         # on the hypotenuse (u+v == det): inclusive
    add(tri0, (0.0, 0.0, 1), (0, 0, -1))         # through vertex 0
    add(tri0, (0.5, 0.0, 1), (0, 0, -1))         # on edge v=0
    add(tri0, (0.0, 0.5, 1), (0, 0, -1))         # on edge u=0
    add(tri0, (-1e-7, 0.25, 1), (0, 0, -1))      # just outside
    add(tri0, (0.25, 0.25, 1e-6), (0, 0, -1))    # t == EPSILON boundary
    add(tri0, (0.25, 0.25, 2e-6), (0, 0, -1))
    add(tri0, (0.25, 0.25, 1), (1, 0, -1e-6))    # det just around EPSILON (un-normalised direction on purpose)
    add(tri0, (0.25, 0.25, 1), (1, 0, -2e-6))
    add(tri0, (0.25, 0.25, 1), (0, 0, 0))        # zero direction
    for _ in range(2000):
        v9 = rng.uniform(-1, 1, 9).astype(np.float32)
        o = rng.uniform(-2, 2, 3).astype(np.float32)
        tgt = (v9[0:3] + v9[3:6] + v9[6:9]) / 3 + rng.normal(0, 0.3, 3).astype(np.float32)
        d = (tgt - o).astype(np.float32)
        d = (d / np.linalg.norm(d)).astype(np.float32)
        add(v9, o, d)
    save("kat_triangle.npz", v9=np.stack([c[0] for c in cases]), o=np.stack([c[1] for c in cases]), d=np.stack([c[2] for c in cases]),
                        hit=np.array([c[3] for c in cases], np.uint8), t=np.array([c[4] for c in cases], np.float32),
                        point=np.stack([c[5] for c in cases]), normal=np.stack([c[6] for c in cases]))
    print("kat_triangle:", len(cases), "cases,", int(sum(c[3] for c in cases)), "hits")

    # ---- identity-transform scenes through the reference's intersectRays
    ident = {}
    cams = {"rocketman": S.camera((0.3, 0.5, 6), fov=60, width=96, height=96),
            "wahoo": S.camera((1, 8, 30), fov=40, width=96, height=96),
            "cube": S.camera((0.2, 0.3, 3), fov=45, width=64, height=64),
            "sphereBlender": S.camera((0, 0, 4), fov=40, width=64, height=64)}
    for name, cam in cams.items():
        sc = ref.scene()
        sc.add_mesh(meshes[name])
        if name == "wahoo":
            sc.add_mesh(meshes["sphere"])  # second geometry, still identity
        rays, _ = ref.camera_rays(cam)
        # plus incoherent rays aimed into the mesh's bounding box
        lo, hi = meshes[name].reshape(-1, 3).min(0), meshes[name].reshape(-1, 3).max(0)
        m = 4096
        o = rng.uniform(lo - 2, hi + 2, (m, 3)).astype(np.float32)
        tgt = rng.uniform(lo, hi, (m, 3)).astype(np.float32)
        d = tgt - o
        d = (d / np.linalg.norm(d, axis=1, keepdims=True)).astype(np.float32)
        rays = np.concatenate([rays, np.concatenate([o, d], 1)]).astype(np.float32)
        r = sc.intersect(rays)
        ident[name + "_rays"] = rays
        for k in ("hit", "geom", "tri", "t", "point", "normal"):
            ident[name + "_" + k] = r[k]
        print("identity", name, "rays", len(rays), "hits", int(r["hit"].sum()))
    save("ref_identity.npz", **ident)

    # ---- config 1 and the main() scene
    for desc, fname in ((S.config1(), "ref_c1.npz"), (S.reference_main(160, 160), "ref_main.npz"), (S.scaled_meshes(160, 160), "ref_scaled.npz")):
        sc = S.instantiate_on(desc, ref.scene(), meshes)
        rays, _ = ref.camera_rays(desc["camera"])
        r = sc.intersect(rays)
        out = dict(rays=rays, **r)
        out["as_coded"] = sc.render_as_coded(desc["camera"])
        save(fname, **out)
        print(desc["name"], "rays", len(rays), "hits", int(r["hit"].sum()), "geoms hit", np.unique(r["geom"]))

    # ---- camera and matrices
    cam_out = {}
    for i, cam in enumerate([S.config1()["camera"], S.camera((1, 2, 3), yaw=-120, pitch=15, fov=55, width=40, height=24),
                             S.config4_camera(32, 18)]):
        rays, basis = ref.camera_rays(cam)
        iv, ip = ref.camera_inverse_matrices(cam)
        cam_out["cam%d" % i] = np.array([*cam["pos"], cam["yaw"], cam["pitch"], cam["fov"], cam["width"], cam["height"],
                                         cam["near"], cam["far"]], np.float64)
        cam_out["rays%d" % i], cam_out["basis%d" % i], cam_out["iv%d" % i], cam_out["ip%d" % i] = rays, basis, iv, ip
    save("ref_camera.npz", **cam_out)
    trs = [((0, 0, 0), (0, 0, 0), (1, 1, 1)), ((0, 0, 0), (0, 90, 180), (1, 1, 1)), ((-0.5, -0.5, 1.0), (0, 0, 0), (1, 1, 1)),
           ((0, 2.5, 0), (90, 0, 0), (5, 5, 5)), ((1.5, -2, 0.25), (33, -71, 12.5), (0.5, 2, 3))]
    mats = []
    for pos, rot, scl in trs:
        sc = ref.scene()
        gi = sc.add_mesh(meshes["triangle"], pos, rot, scl)
        m, inv = sc.matrices(gi)
        mats.append(np.concatenate([m, inv]))
    save("ref_matrices.npz", trs=np.array(trs, np.float32), mats=np.stack(mats))
    total = sum(os.path.getsize(os.path.join(OUT, f)) for f in os.listdir(OUT))
    print("golden bytes:", total)


if __name__ == "__main__":
    main()
