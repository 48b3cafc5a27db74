"""CPU: the oracle restatement (oracle/firefly_oracle.cpp) against the golden vectors that tools/make_golden.py
produced from the reference's own routine text (oracle/_ref), and -- where that library is present -- against it live.
This is what pins the oracle; the GPU parity tests then compare the CUDA path with the oracle."""
import numpy as np
import pytest

from conftest import assert_hits_equal, bits
from gpupathtracer_b200 import scenes as S


def test_triangle_known_answers(orc, golden):
    g = golden("kat_triangle.npz")
    n = len(g["hit"])
    assert n > 2000 and g["hit"].sum() > 100
    for i in range(n):
        hit, t, p, nn = orc.intersect_triangle(g["v9"][i], g["o"][i], g["d"][i])
        assert hit == bool(g["hit"][i]), i
        if hit:
            assert bits(t) == bits(g["t"][i]), i
            assert (bits(p) == bits(g["point"][i])).all(), i
            assert (bits(nn) == bits(g["normal"][i])).all(), i
    # analytic pins: front hit at t=1, back face culled, inclusive edges
    assert g["hit"][0] == 1 and g["t"][0] == 1.0
    assert g["hit"][1] == 0
    assert g["hit"][2] == 1 and g["hit"][3] == 1 and g["hit"][4] == 1 and g["hit"][5] == 1
    assert g["hit"][6] == 0


@pytest.mark.parametrize("name", ["rocketman", "wahoo", "cube", "sphereBlender"])
def test_identity_scenes_bit_exact(orc, golden, meshes, name):
    g = golden("ref_identity.npz")
    sc = orc.scene()
    sc.add_mesh(meshes[name])
    if name == "wahoo":
        sc.add_mesh(meshes["sphere"])
    want = {k: g[name + "_" + k] for k in ("hit", "geom", "tri", "t", "point", "normal")}
    for mode in (0, 1):  # the reference's object-space path, and the baked path the GPU implements
        got = sc.intersect(g[name + "_rays"], mode=mode)
        assert_hits_equal(got, want, "%s mode %d" % (name, mode))
        assert (got["hit"] == want["hit"]).all()
        h = want["hit"].astype(bool)
        assert (bits(got["point"][h]) == bits(want["point"][h])).all()
        if mode == 0:
            assert (bits(got["normal"][h]) == bits(want["normal"][h])).all()
        else:  # baked mode skips the multiply by inverse(transpose(I)), which turns -0 into +0: same values
            assert (got["normal"][h] == want["normal"][h]).all()


def test_cube_semantics(orc, meshes):
    """Back-face culling pin: from outside exactly the front faces are visible, from inside nothing."""
    sc = orc.scene()
    sc.add_mesh(meshes["cube"])
    rng = np.random.default_rng(1)
    d = rng.normal(size=(512, 3)).astype(np.float32)
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    inside = np.concatenate([np.zeros((512, 3), np.float32), d], 1)
    assert sc.intersect(inside)["hit"].sum() == 0
    outside = np.concatenate([-3 * d, d], 1).astype(np.float32)
    r = sc.intersect(outside)
    assert r["hit"].all()
    assert np.allclose(r["t"], np.linalg.norm(r["point"] + 3 * d, axis=1), rtol=1e-5)


def test_config1_reference_path(orc, golden, meshes):
    """Config 1 has a translated triangle: mode 0 (object-space, the reference's own path) must be bit-exact; the baked
    mode the GPU uses must give the same ids and t within 1e-5 relative (BASELINE.json north_star)."""
    g = golden("ref_c1.npz")
    desc = S.config1()
    sc = S.instantiate_on(desc, orc.scene(), meshes)
    rays, _ = orc.camera_rays(desc["camera"])
    assert (bits(rays) == bits(g["rays"])).all()
    got0 = sc.intersect(rays, mode=0)
    assert_hits_equal(got0, g, "C1 mode 0")
    got1 = sc.intersect(rays, mode=1)
    assert (got1["geom"] == g["geom"]).all() and (got1["tri"] == g["tri"]).all()
    h = g["hit"].astype(bool)
    assert np.allclose(got1["t"][h], g["t"][h], rtol=1e-5, atol=0)
    assert (sc.render_as_coded(desc["camera"], mode=0) == g["as_coded"]).all()


def test_main_scene_with_planes(orc, golden, meshes):
    g = golden("ref_main.npz")
    desc = S.reference_main(160, 160)
    sc = S.instantiate_on(desc, orc.scene(), meshes)
    got = sc.intersect(g["rays"], mode=0)
    assert_hits_equal(got, g, "main() scene mode 0")
    assert (sc.render_as_coded(desc["camera"], mode=0) == g["as_coded"]).all()
    got1 = sc.intersect(g["rays"], mode=1)
    assert (got1["geom"] == g["geom"]).all() and (got1["tri"] == g["tri"]).all()
    h = g["hit"].astype(bool)
    assert np.allclose(got1["t"][h], g["t"][h], rtol=1e-5)


def test_scaled_meshes_as_coded_frame(orc, golden, meshes):
    """Meshes under scales below / above 1 (scenes.scaled_meshes): the reference's frame shows inverse(transpose(M)) * n
    un-normalised (kernel.cu:117,183) and make_uchar3 wraps above 255 (kernel.cu:214).  Mode 0 is the reference's own path
    (bit-exact, frame included); mode 1 -- the GPU contract: hit found on baked triangles, same colour rule -- must show the
    same bytes wherever it finds the same primitive."""
    g = golden("ref_scaled.npz")
    desc = S.scaled_meshes(160, 160)
    sc = S.instantiate_on(desc, orc.scene(), meshes)
    got = sc.intersect(g["rays"], mode=0)
    assert_hits_equal(got, g, "scaled scene mode 0")
    frame = g["as_coded"]
    assert (sc.render_as_coded(desc["camera"], mode=0) == frame).all()
    assert frame.max() < 255 and (frame.reshape(-1, 3)[g["geom"] == 0].astype(int).sum(-1) > 300).any()  # |n| = 2 on the half-scale mesh wraps: no clamped 255s, sums a unit normal cannot reach
    got1 = sc.intersect(g["rays"], mode=1)
    same = (got1["geom"] == g["geom"]) & (got1["tri"] == g["tri"])
    assert same.mean() > 0.999
    f1 = sc.render_as_coded(desc["camera"], mode=1).reshape(-1, 3)
    assert (f1[same] == frame.reshape(-1, 3)[same]).all()


def test_camera_and_matrices(orc, golden):
    g = golden("ref_camera.npz")
    for i in range(3):
        c = g["cam%d" % i]
        cam = S.camera(c[0:3], yaw=c[3], pitch=c[4], fov=c[5], width=int(c[6]), height=int(c[7]), near=c[8], far=c[9])
        rays, basis = orc.camera_rays(cam)
        iv, ip = orc.camera_inverse_matrices(cam)
        assert (bits(rays) == bits(g["rays%d" % i])).all()
        assert (bits(basis) == bits(g["basis%d" % i])).all()
        assert (bits(iv) == bits(g["iv%d" % i])).all() and (bits(ip) == bits(g["ip%d" % i])).all()
    m = golden("ref_matrices.npz")
    tri = np.array([[0, 0, 0, 1, 0, 0, 0, 1, 0]], np.float32)
    for trs, want in zip(m["trs"], m["mats"]):
        sc = orc.scene()
        gi = sc.add_mesh(tri, trs[0], trs[1], trs[2])
        mm, inv = sc.matrices(gi)
        assert (bits(np.concatenate([mm, inv])) == bits(want)).all()


def test_philox_known_answers(orc):
    """Random123 kat_vectors for philox4x32_10."""
    kats = [((0, 0, 0, 0), (0, 0), (0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8)),
            ((0xffffffff,) * 4, (0xffffffff,) * 2, (0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd)),
            ((0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344), (0xa4093822, 0x299f31d0), (0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1))]
    for ctr, key, want in kats:
        assert tuple(int(x) for x in orc.philox(ctr, key)) == want


def test_sampling_primitives(orc):
    u = np.linspace(0, 1, 4097, dtype=np.float32)[:-1]
    s, c = orc.sincos2pi(u)
    assert np.abs(s - np.sin(2 * np.pi * u.astype(np.float64))).max() < 2e-7
    assert np.abs(c - np.cos(2 * np.pi * u.astype(np.float64))).max() < 2e-7
    rng = np.random.default_rng(3)
    for _ in range(200):
        n = rng.normal(size=3)
        n = (n / np.linalg.norm(n)).astype(np.float32)
        t, b = orc.onb(n)
        assert abs(np.dot(t, b)) < 1e-6 and abs(np.dot(t, n)) < 1e-6 and abs(np.dot(b, n)) < 1e-6
        assert abs(np.linalg.norm(t) - 1) < 1e-6 and abs(np.linalg.norm(b) - 1) < 1e-6
    w = orc.cosine_sample_hemisphere(0.25, 0.125)
    assert abs(np.linalg.norm(w) - 1) < 1e-6 and w[2] > 0


def test_oracle_vs_reference_live(orc, ref, meshes):
    """Where oracle/_ref is present: fresh random rays, restatement vs the reference's own text, bit for bit."""
    rng = np.random.default_rng(99)
    desc = S.reference_main(64, 64)
    a = S.instantiate_on(desc, orc.scene(), meshes)
    b = S.instantiate_on(desc, ref.scene(), meshes)
    o = rng.uniform(-6, 6, (3000, 3)).astype(np.float32)
    d = rng.normal(size=(3000, 3)).astype(np.float32)
    rays = np.concatenate([o, d], 1)  # un-normalised directions on purpose
    assert_hits_equal(a.intersect(rays, mode=0), b.intersect(rays), "live main()")
    assert (a.render_as_coded(desc["camera"], mode=0) == b.render_as_coded(desc["camera"])).all()


def test_path_tracer_properties(orc, meshes):
    """The as-intended tracer (no reference counterpart): furnace test and determinism."""
    cube = S.flip_winding(meshes["cube"])
    sc = orc.scene()
    sc.add_mesh(cube, scale=(4, 4, 4), material=dict(type=1, albedo=(1, 1, 1)))
    cam = S.camera((0, 0, 0), fov=60, width=8, height=8)
    img, n = sc.render(cam, 4, 3, background=(1, 1, 1))
    assert (img[..., :3] == 0).all()  # closed box, no light, paths die at max bounce without contribution
    sc2 = orc.scene()
    sc2.add_mesh(cube, scale=(4, 4, 4), material=dict(type=0, emissive=(1, 0.5, 0.25), intensity=2.0))
    img2, n2 = sc2.render(cam, 2, 3)
    assert np.allclose(img2[..., :3], [2, 1, 0.5]) and n2 == 8 * 8 * 2
    a, _ = sc.render(cam, 4, 3, seed=7)
    b, _ = sc.render(cam, 2, 3, seed=7)
    c, _ = sc.render(cam, 2, 3, seed=7, first_sample=2)
    assert np.allclose(a, (b + c) / 2, atol=1e-6)


def test_next_event_estimation_agrees_with_plain_loop(orc, meshes):
    """NEE (shadow rays to the emitter triangles) and the plain loop estimate the same image: on the C2 box the two means
    agree within Monte-Carlo noise, NEE with far less variance; with no mesh emitter NEE changes nothing at all."""
    desc = S.config2(meshes, 24, 24, 1, 3)
    sc = S.instantiate_on(desc, orc.scene(), meshes)
    cam = desc["camera"]
    plain, n0 = sc.render(cam, 192, 3, seed=5)
    nee, n1, ns = sc.render(cam, 48, 3, seed=6, nee=1, want_shadow_rays=True)
    assert ns > 0 and n1 > 0
    a, b = plain[..., :3].mean(), nee[..., :3].mean()
    assert abs(a - b) <= 0.04 * a, (a, b)  # whole-image means: within 4 %
    # per-pixel: the smoothed difference stays small (both are unbiased estimates of the same frame)
    blk = lambda im: im[..., :3].reshape(6, 4, 6, 4, 3).mean(axis=(1, 3))
    assert np.abs(blk(plain) - blk(nee)).max() <= 0.25 * max(1e-3, blk(plain).max())
    # determinism and sample additivity hold with NEE too
    x, _ = sc.render(cam, 4, 3, seed=7, nee=1)
    y, _ = sc.render(cam, 2, 3, seed=7, nee=1)
    z, _ = sc.render(cam, 2, 3, seed=7, nee=1, first_sample=2)
    assert np.allclose(x, (y + z) / 2, atol=1e-6)
    # a scene whose only emitter is missing: nee=1 is bit-identical to nee=0
    d1 = S.config1(32, 32)
    s1 = S.instantiate_on(d1, orc.scene(), meshes)
    p0, _ = s1.render(d1["camera"], 2, 2, background=(0.2, 0.3, 0.4))
    p1, _ = s1.render(d1["camera"], 2, 2, background=(0.2, 0.3, 0.4), nee=1)
    assert (p0.view(np.uint32) == p1.view(np.uint32)).all()


def test_sphere_primitive_known_answers(orc):
    """GeometryType::SPHERE (declared by the reference, branch left empty: kernel.cu:166-169): analytic known answers of
    the restatement, and the object-space path (mode 0) against the baked path (mode 1)."""
    sc = orc.scene()
    sc.add_sphere(1.0)
    rays = np.float32([[0, 0, 5, 0, 0, -1],      # head on: enters at z = 1, t = 4
                       [0, 0, 0, 0, 0, -1],      # starts inside: front side only -> miss
                       [1, 0, 5, 0, 0, -1],      # tangent: discriminant 0 -> miss
                       [0.6, 0, 5, 0, 0, -2],    # un-normalised direction; enters at z = 0.8
                       [0, 0, -5, 0, 0, -1]])    # pointing away
    for mode in (0, 1):
        h = sc.intersect(rays, mode=mode)
        assert h["hit"].tolist() == [1, 0, 0, 1, 0]
        assert h["t"][0] == 4.0 and np.allclose(h["point"][0], [0, 0, 1]) and np.allclose(h["normal"][0], [0, 0, 1])
        assert np.isclose(h["t"][3], 4.2, atol=1e-6) and h["geom"][0] == 0 and h["tri"][0] == -1
    # transformed: radius 0.5, scaled (2, 1, 1), moved to x = 3 -> world semi-axes (1, .5, .5)
    sc2 = orc.scene()
    sc2.add_sphere(0.5, pos=(3, 0, 0), scale=(2, 1, 1))
    r2 = np.float32([[3, 0, 5, 0, 0, -1], [10, 0, 0, -1, 0, 0], [3, 0.6, 5, 0, 0, -1]])
    a, b = sc2.intersect(r2, mode=0), sc2.intersect(r2, mode=1)
    assert a["hit"].tolist() == [1, 1, 0] and (a["t"].view(np.uint32) == b["t"].view(np.uint32)).all()
    assert np.isclose(a["t"][0], 4.5) and np.isclose(a["t"][1], 6.0)
    assert np.allclose(a["normal"][1] / np.linalg.norm(a["normal"][1]), [1, 0, 0])


def test_mirror_and_glass_events(orc, meshes):
    """MIRROR / GLASS (BXDFTyp utilities.h:72-73): a mirror sphere under a uniform sky returns the sky tinted by its
    specular colour after one bounce; glass splits between the two lobes with Schlick's Fresnel."""
    sky = (0.5, 0.75, 1.0)
    cam = S.camera((0, 0, 4), fov=10, width=6, height=6)
    m = orc.scene()
    m.add_sphere(1.0, material=dict(type=2, specular=(0.5, 1.0, 0.25)))
    img, n = m.render(cam, 3, 2, background=sky)
    assert np.allclose(img[..., :3], np.float32(sky) * np.float32([0.5, 1.0, 0.25]), atol=1e-6) and n == 6 * 6 * 3 * 2
    g = orc.scene()
    g.add_sphere(1.0, material=dict(type=3, ior=1.5, specular=(1, 0, 0), transmittance=(0, 0, 1)))
    big = S.camera((0, 0, 4), fov=2, width=32, height=32)  # near-normal incidence: F = r0 = 0.04
    img, _ = g.render(big, 64, 1, background=(1, 1, 1), aa=1)
    red, blue = img[..., 0].mean(), img[..., 2].mean()
    assert abs(red - 0.04) < 0.01 and abs(blue - 0.96) < 0.01 and np.allclose(img[..., 1], 0)
    # refraction bends towards the normal: a ray entering off-axis leaves the front interface closer to -n
    off = S.camera((0.9, 0, 4), yaw=-90, fov=0.01, width=1, height=1)
    o = g.render(off, 1, 0, background=(0, 0, 0))  # max_bounces 0: nothing to add, just must not crash
    assert o[0].shape == (1, 1, 4)


def test_shading_attribute_known_answers(orc):
    """Barycentrics (kernel.cu:80-81), interpolated normal and uv on one triangle, against hand-computed values."""
    sc = orc.scene()
    tri = np.array([[0, 0, 0, 1, 0, 0, 0, 1, 0]], np.float32)
    gi = sc.add_mesh(tri)
    sc.set_normals(gi, np.array([[0, 0, 1, 1, 0, 1, 0, 1, 1]], np.float32))
    sc.set_uvs(gi, np.array([[0, 0, 1, 0, 0, 1]], np.float32))
    rays = np.array([[0.25, 0.5, 1, 0, 0, -1], [0.125, 0.125, 2, 0, 0, -3], [2, 2, 1, 0, 0, -1]], np.float32)
    h = sc.intersect_uv(rays)
    assert h["geom"].tolist() == [0, 0, -1] and h["tri"].tolist() == [0, 0, -1]
    assert h["u"].tolist() == [0.25, 0.125, 0.0] and h["v"].tolist() == [0.5, 0.125, 0.0]
    cam = dict(pos=(0.25, 0.5, 1.0), world_up=(0, 1, 0), yaw=-90.0, pitch=0.0, fov=40.0, width=2, height=2, near=0.1, far=100.0)
    # pixel (1,1) of a 2x2 frame is the image centre (integer pixel corners, kernel.cu:200-201): the ray straight down -z
    n = sc.render_aov(cam, 3)[1, 1]
    want = np.array([0.25, 0.5, 1.0]) / np.linalg.norm([0.25, 0.5, 1.0])
    assert np.allclose(n[:3], want, atol=1e-6) and n[3] == 1.0
    assert np.allclose(sc.render_aov(cam, 4)[1, 1], [0.25, 0.5, 0, 1], atol=1e-6)
    assert np.allclose(sc.render_aov(cam, 5)[1, 1], [0.25, 0.25, 0.5, 1], atol=1e-6)
    assert np.allclose(sc.render_aov(cam, 6)[1, 1], [0.8, 0.8, 0.8, 1])
    sc.set_normals(gi, None)
    assert np.allclose(sc.render_aov(cam, 3)[1, 1], [0, 0, 1, 1])   # geometric normal, facing the camera


def test_shading_attribute_properties(orc, meshes):
    """All-zero normals and all -1 face materials change nothing; per-face materials equal the same triangles split into
    one geometry per material (same flat order, so the same light list and the same tie-breaks); smooth normals change
    shading but not what is hit."""
    attrs = S.load_mesh_attrs()
    desc = S.shaded(meshes, 40, 32, 3, 4)
    cam = desc["camera"]
    base = S.instantiate_on(desc, orc.scene(), meshes, attrs)
    img, rays = base.render(cam, 3, 4, seed=5, aa=1, nee=1)
    assert img[..., :3].max() > 0 and np.isfinite(img).all()
    # (1) no-op attributes
    plain_desc = S.config2(meshes, 40, 32, 3, 4)
    plain = S.instantiate_on(plain_desc, orc.scene(), meshes)
    want, _ = plain.render(cam, 3, 4, seed=5, aa=1, nee=1)
    noop = S.instantiate_on(plain_desc, orc.scene(), meshes)
    nr = len(meshes["rocketman"])
    noop.set_normals(2, np.zeros((nr, 9), np.float32))
    noop.set_face_materials(2, np.full(nr, -1, np.int32), [dict(type=0, emissive=(9, 9, 9), intensity=9.0)])
    got, _ = noop.render(cam, 3, 4, seed=5, aa=1, nee=1)
    assert (bits(got) == bits(want)).all()
    # (2) per-face materials == split geometries
    mats = [dict(type=1, albedo=(0.9, 0.2, 0.2)), dict(type=0, emissive=(1, 0.8, 0.6), intensity=4.0), dict(type=2, specular=(0.9, 0.9, 0.9))]
    w = meshes["wahoo"]
    cut = [0, 1700, 3400, len(w)]
    a, b = orc.scene(), orc.scene()
    for s in (a, b):
        s.add_mesh(S.flip_winding(meshes["cube"]), scale=(5, 5, 5))
    gi = a.add_mesh(w, pos=(0, -2.4, 0), scale=(0.2, 0.2, 0.2))
    fm = np.concatenate([np.full(cut[k + 1] - cut[k], k, np.int32) for k in range(3)])
    a.set_face_materials(gi, fm, mats)
    for k in range(3):
        b.add_mesh(w[cut[k]:cut[k + 1]], pos=(0, -2.4, 0), scale=(0.2, 0.2, 0.2), material=mats[k])
    for nee in (0, 1):
        ia, ra, sa = a.render(cam, 3, 4, seed=9, aa=1, nee=nee, want_shadow_rays=True)
        ib, rb, sb = b.render(cam, 3, 4, seed=9, aa=1, nee=nee, want_shadow_rays=True)
        assert (bits(ia) == bits(ib)).all() and ra == rb and sa == sb and ia[..., :3].max() > 0
    # (3) smooth normals: same primary hits, different shading normals on the curved meshes
    flat_desc = dict(desc, geoms=[dict(g, normals=None) for g in desc["geoms"]])
    flat = S.instantiate_on(flat_desc, orc.scene(), meshes, attrs)
    rays6, _ = orc.camera_rays(cam)
    h1, h2 = base.intersect(rays6, mode=1), flat.intersect(rays6, mode=1)
    assert (h1["geom"] == h2["geom"]).all() and (h1["tri"] == h2["tri"]).all() and (bits(h1["t"]) == bits(h2["t"])).all()
    n1, n2 = base.render_aov(cam, 3), flat.render_aov(cam, 3)
    on_sphere = np.isin(h1["geom"].reshape(32, 40), [2, 3, 4, 5])   # wahoo.obj and sphere.obj have smooth normals
    assert on_sphere.sum() > 50
    diff = np.abs(n1 - n2)[..., :3].max(-1)
    assert (diff[on_sphere] > 1e-3).mean() > 0.8 and (diff[~on_sphere & (h1["geom"].reshape(32, 40) < 2)] == 0).all()
    assert np.allclose(np.linalg.norm(n1[..., :3], axis=-1)[h1["geom"].reshape(32, 40) >= 0], 1.0, atol=1e-5)
    img_flat, _ = flat.render(cam, 3, 4, seed=5, aa=1, nee=1)
    assert not (bits(img_flat) == bits(img)).all()
