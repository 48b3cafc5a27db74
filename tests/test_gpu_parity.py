"""GPU: closest-hit parity of the CUDA path (through the C ABI) against the CPU oracle -- bit-exact geometryIndex,
triangleIndex and key t -- on the reference's scenes, on edge cases, and on synthetic soups (oracle on a ray subset,
GPU brute force as the bridge at full size)."""
import numpy as np
import pytest

from conftest import assert_hits_equal, bits
from gpupathtracer_b200 import scenes as S

pytestmark = pytest.mark.gpu
MODES = ["BVH8", "BVH2", "BRUTE"]


def mode_of(ff, name):
    return {"BVH8": ff.TRACE_BVH8, "BVH2": ff.TRACE_BVH2, "BRUTE": ff.TRACE_BRUTE}[name]


def random_rays(rng, lo, hi, n, normalise=True):
    o = rng.uniform(lo - 0.5 * (hi - lo) - 0.1, hi + 0.5 * (hi - lo) + 0.1, (n, 3)).astype(np.float32)
    tgt = rng.uniform(lo, hi, (n, 3)).astype(np.float32)
    d = (tgt - o).astype(np.float32)
    if normalise:
        d = (d / np.maximum(np.linalg.norm(d, axis=1, keepdims=True), 1e-12)).astype(np.float32)
    return np.concatenate([o, d], 1).astype(np.float32)


@pytest.mark.parametrize("mode", MODES)
def test_golden_identity_scenes(ff, ctx, golden, meshes, mode):
    """Straight against the reference's own outputs (tests/golden/ref_identity.npz)."""
    g = golden("ref_identity.npz")
    for name in ("rocketman", "wahoo", "cube", "sphereBlender"):
        sc = ctx.scene()
        sc.add_triangles(meshes[name])
        if name == "wahoo":
            sc.add_triangles(meshes["sphere"])
        sc.commit()
        got = sc.trace_closest(g[name + "_rays"], mode_of(ff, mode))
        want = {k: g[name + "_" + k] for k in ("geom", "tri", "t")}
        assert_hits_equal(got, want, "%s %s" % (name, mode))


@pytest.mark.parametrize("mode", MODES)
@pytest.mark.parametrize("cfg", ["C1", "C2", "C3", "main"])
def test_config_scenes_vs_oracle(ff, ctx, orc, meshes, golden, cfg, mode):
    desc = {"C1": S.config1(), "C2": S.config2(meshes, 160, 160), "C3": S.config3(meshes, 192, 108),
            "main": S.reference_main(128, 128)}[cfg]
    sc, cam = S.instantiate(desc, ctx, meshes)
    osc = S.instantiate_on(desc, orc.scene(), meshes)
    rays = ctx.camera_rays(cam)
    orays, _ = orc.camera_rays(desc["camera"])
    assert (bits(rays) == bits(orays)).all()
    flat, fg, ft = sc.flat()
    oflat, og, ot = osc.flat()
    assert (bits(flat) == bits(oflat)).all() and (fg == og).all() and (ft == ot).all()
    lo, hi = flat.reshape(-1, 3).min(0), flat.reshape(-1, 3).max(0)
    rng = np.random.default_rng({"C1": 1, "C2": 2, "C3": 3, "main": 4}[cfg])
    extra = np.concatenate([random_rays(rng, lo, hi, 20000), random_rays(rng, lo, hi, 2000, normalise=False)])
    allrays = np.concatenate([orays, extra])
    want = osc.intersect(allrays, mode=1)
    got = sc.trace_closest(allrays, mode_of(ff, mode))
    assert_hits_equal(got, want, "%s %s" % (cfg, mode))
    assert want["hit"].sum() > 1000
    if cfg == "C1":  # and against the reference's own object-space path: same ids, t within 1e-5 relative
        g = golden("ref_c1.npz")
        n = len(g["rays"])
        assert (got["geom"][:n] == g["geom"]).all() and (got["tri"][:n] == g["tri"]).all()
        h = g["hit"].astype(bool)
        assert np.allclose(got["t"][:n][h], g["t"][h], rtol=1e-5, atol=0)


@pytest.mark.parametrize("mode", MODES)
def test_ties_lowest_index_wins(ff, ctx, orc, meshes, mode):
    """The same mesh twice, and shared edges / vertices: strict '<' in scan order means the lowest (geometry, triangle) wins."""
    sc, osc = ctx.scene(), orc.scene()
    for s in (sc, osc):
        (s.add_triangles if s is sc else s.add_mesh)(meshes["sphereBlender"])
        (s.add_triangles if s is sc else s.add_mesh)(meshes["cube"])
        (s.add_triangles if s is sc else s.add_mesh)(meshes["sphereBlender"])
        (s.add_triangles if s is sc else s.add_mesh)(meshes["cube"])
    sc.commit()
    rng = np.random.default_rng(4)
    rays = random_rays(rng, np.float32([-1, -1, -1]), np.float32([1, 1, 1]), 30000)
    # rays aimed exactly at mesh vertices and edge midpoints
    v = meshes["sphereBlender"].reshape(-1, 3, 3)
    targets = np.concatenate([v[:, 0], 0.5 * (v[:, 0] + v[:, 1]), meshes["cube"].reshape(-1, 3, 3)[:, 0],
                              0.5 * (meshes["cube"].reshape(-1, 3, 3)[:, 0] + meshes["cube"].reshape(-1, 3, 3)[:, 1])]).astype(np.float32)
    o = (targets * 3 + np.float32([0.01, 0.02, 0.03])).astype(np.float32)
    d = (targets - o).astype(np.float32)
    rays = np.concatenate([rays, np.concatenate([o, d], 1)])
    want = osc.intersect(rays, mode=1)
    got = sc.trace_closest(rays, mode_of(ff, mode))
    assert_hits_equal(got, want, "ties " + mode)
    assert set(np.unique(got["geom"])) <= {-1, 0, 1}  # duplicates (geometries 2, 3) never win


@pytest.mark.parametrize("mode", MODES)
def test_edge_cases(ff, ctx, orc, meshes, mode):
    m = mode_of(ff, mode)
    rng = np.random.default_rng(8)
    rays = random_rays(rng, np.float32([-1, -1, -1]), np.float32([1, 1, 1]), 4096)
    axis = np.zeros((6, 6), np.float32)  # axis-parallel rays with exact zeros in the direction
    for k in range(6):
        axis[k, k % 3] = -3 if k < 3 else 3
        axis[k, 3 + k % 3] = 1 if k < 3 else -1
    rays = np.concatenate([rays, axis, np.float32([[0, 0, 0, 0, 0, 0]]), np.float32([[np.nan, 0, 0, 1, 0, 0]])])
    # empty scene, and empty batch
    sc = ctx.scene()
    sc.commit()
    r = sc.trace_closest(rays, m)
    assert (r["geom"] == -1).all() and np.isinf(r["t"]).all()
    assert len(sc.trace_closest(np.zeros((0, 6), np.float32), m)["geom"]) == 0
    # 1..9 triangles (tiny trees), degenerate triangles mixed in
    for n in (1, 2, 3, 4, 5, 9):
        tris = meshes["cube"][:n].copy()
        if n >= 4:
            tris[1, 3:6] = tris[1, 0:3]  # zero-area triangle
        sc, osc = ctx.scene(), orc.scene()
        sc.add_triangles(tris)
        osc.add_mesh(tris)
        sc.commit()
        assert_hits_equal(sc.trace_closest(rays, m), osc.intersect(rays, mode=1), "n=%d %s" % (n, mode))
    # all triangles identical (Morton duplicates), and a mesh far from the origin with a large extent
    same = np.repeat(meshes["triangle"], 300, 0)
    big = (meshes["sphereBlender"] * 500 + 2000).astype(np.float32)
    for tris, lo, hi in ((same, np.float32([0, 0, -1]), np.float32([1, 1, 1])), (big, np.float32([1500] * 3), np.float32([2500] * 3))):
        sc, osc = ctx.scene(), orc.scene()
        sc.add_triangles(tris)
        osc.add_mesh(tris)
        sc.commit()
        rr = random_rays(rng, lo, hi, 8192)
        want = osc.intersect(rr, mode=1)
        assert_hits_equal(sc.trace_closest(rr, m), want, mode)
        assert want["hit"].sum() > 100


def test_transforms_match_oracle(ff, ctx, orc, meshes, golden):
    mats = golden("ref_matrices.npz")
    sc, osc = ctx.scene(), orc.scene()
    for trs in mats["trs"]:
        sc.add_triangles(meshes["sphere"], trs[0], trs[1], trs[2])
        osc.add_mesh(meshes["sphere"], trs[0], trs[1], trs[2])
    sc.commit()
    for gi, want in enumerate(mats["mats"]):
        m, inv = sc.matrices(gi)
        assert (bits(np.concatenate([m, inv])) == bits(want)).all()
    flat, _, _ = sc.flat()
    oflat, _, _ = osc.flat()
    assert (bits(flat) == bits(oflat)).all()
    rng = np.random.default_rng(12)
    rays = random_rays(rng, flat.reshape(-1, 3).min(0), flat.reshape(-1, 3).max(0), 20000)
    assert_hits_equal(sc.trace_closest(rays), osc.intersect(rays, mode=1), "transformed spheres")


@pytest.mark.parametrize("n_tris", [20000, 200000])
def test_soup_vs_oracle_and_bridge(ff, ctx, orc, n_tris):
    """Config-4 generator at reduced size: generators bit-identical to the oracle's, BVH == oracle on a subset,
    BVH == GPU brute force on a larger batch."""
    soup = ctx.make_soup(n_tris)
    assert (bits(soup) == bits(orc.make_soup(n_tris))).all()
    rays = ctx.make_rays_uniform(0, 1 << 17)
    assert (bits(rays[:8192]) == bits(orc.make_rays_uniform(0, 8192))).all()
    assert (bits(ctx.make_rays_uniform(5000, 100, salt=3)) == bits(orc.make_rays_uniform(5000, 100, salt=3))).all()
    sc, osc = ctx.scene(), orc.scene()
    sc.add_triangles(soup)
    osc.add_mesh(soup)
    sc.commit()
    info = sc.info()
    assert info["n_triangles"] == n_tris and info["n_nodes"] > 0 and info["depth"] <= 64
    sub = rays[:4096 if n_tris <= 20000 else 1024]
    want = osc.intersect(sub, mode=1)
    for mode in (ff.TRACE_BVH8, ff.TRACE_BVH2, ff.TRACE_BRUTE):
        assert_hits_equal(sc.trace_closest(sub, mode), want, "soup subset")
    nb = 1 << 17 if n_tris <= 20000 else 1 << 15
    brute = sc.trace_closest(rays[:nb], ff.TRACE_BRUTE)
    assert_hits_equal(sc.trace_closest(rays[:nb], ff.TRACE_BVH8), brute, "bvh8 vs brute")
    assert_hits_equal(sc.trace_closest(rays[:nb], ff.TRACE_BVH2), brute, "bvh2 vs brute")
    assert 0.5 < brute["hit"].mean() <= 1.0


def test_full_size_config4_properties(ff, ctx):
    """BASELINE config 4 at full size (1M triangles): properties that need no oracle -- the two hierarchies agree on
    every ray, results are independent of batch order, and a hit's t reproduces from (ray, triangle)."""
    n = 1000000
    soup = ctx.make_soup(n)
    sc = ctx.scene()
    sc.add_triangles(soup)
    sc.commit()
    nr = 1 << 20
    rays = ctx.make_rays_uniform(0, nr)
    a = sc.trace_closest(rays, ff.TRACE_BVH8)
    b = sc.trace_closest(rays, ff.TRACE_BVH2)
    assert_hits_equal(a, b, "bvh8 vs bvh2 at 1M")
    perm = np.random.default_rng(0).permutation(nr)
    c = sc.trace_closest(rays[perm], ff.TRACE_BVH8)
    assert (c["tri"] == a["tri"][perm]).all() and (bits(c["t"]) == bits(a["t"][perm])).all()
    brute = sc.trace_closest(rays[:4096], ff.TRACE_BRUTE)
    assert_hits_equal({k: v[:4096] for k, v in a.items()}, brute, "bvh8 vs brute at 1M")
    h = a["geom"] >= 0
    assert 0.9 < h.mean()
    # geometric check of t in float64 against the hit triangle's plane
    idx = np.nonzero(h)[0][:50000]
    tri = soup[a["tri"][idx]].astype(np.float64).reshape(-1, 3, 3)
    nrm = np.cross(tri[:, 1] - tri[:, 0], tri[:, 2] - tri[:, 0])
    o, d = rays[idx, :3].astype(np.float64), rays[idx, 3:].astype(np.float64)
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    t = np.einsum("ij,ij->i", tri[:, 0] - o, nrm) / np.einsum("ij,ij->i", d, nrm)
    assert np.abs(t - a["t"][idx]).max() < 1e-4


@pytest.mark.parametrize("mode", MODES)
def test_any_hit_consistent_with_closest(ff, ctx, meshes, mode):
    desc = S.config2(meshes, 64, 64)
    sc, cam = S.instantiate(desc, ctx, meshes)
    rng = np.random.default_rng(6)
    rays = random_rays(rng, np.float32([-2.5] * 3), np.float32([2.5] * 3), 20000)
    m = mode_of(ff, mode)
    closest = sc.trace_closest(rays, m)
    assert (sc.trace_any(rays, None, m) == (closest["geom"] >= 0)).all()
    tmax = rng.uniform(0.1, 6, len(rays)).astype(np.float32)
    occ = sc.trace_any(rays, tmax, m)
    assert (occ == (closest["t"] < tmax)).all()


def test_analytic_only_scene_and_plane_occluders_finish_every_ray(ff, ctx, orc, meshes):
    """Rays that end in the fetch path of the persistent kernels (a scene without any triangle hierarchy; shadow rays
    stopped by an analytic plane before traversal starts) must not end a warp's loop early: batches far larger than the
    resident grid, every ray checked."""
    n = 600000
    rng = np.random.default_rng(12)
    rays = np.empty((n, 6), np.float32)
    rays[:, :3] = rng.uniform((-1.5, -1.5, 2.0), (1.5, 1.5, 6.0), (n, 3))
    rays[:, 3:] = np.float32((0, 0, -1)) + rng.uniform(-0.05, 0.05, (n, 3)).astype(np.float32)
    # planes only: no nodes at all
    sc = ctx.scene()
    sc.add_plane((0, 0, 0), (0, 0, 0), (4, 4, 4))
    sc.add_plane((0, 0, -1), (0, 0, 0), (8, 8, 8))
    sc.commit()
    osc = orc.scene()
    osc.add_plane((0, 0, 0), (0, 0, 0), (4, 4, 4))
    osc.add_plane((0, 0, -1), (0, 0, 0), (8, 8, 8))
    want = osc.intersect(rays[:50000])
    got = sc.trace_closest(rays)
    assert (got["geom"][:50000] == want["geom"]).all() and (bits(got["t"][:50000]) == bits(want["t"])).all()
    assert (got["geom"] >= 0).mean() > 0.9 and (got["geom"] == sc.trace_closest(rays, ff.TRACE_BVH8 | ff.TRACE_STATIC)["geom"]).all()
    assert (sc.trace_any(rays) == (got["geom"] >= 0)).all()
    # a mesh hidden behind a plane: every any-hit query is decided by the plane, in the fetch path
    sc2 = ctx.scene()
    sc2.add_plane((0, 0, 1), (0, 0, 0), (20, 20, 20))
    sc2.add_triangles(meshes["cube"])
    sc2.commit()
    occ = sc2.trace_any(rays)
    assert occ.all()
    c2 = sc2.trace_closest(rays)
    assert (c2["geom"] == 0).all()


def test_instrumented_counts_and_device_entry(ff, ctx):
    soup = ctx.make_soup(50000)
    sc = ctx.scene()
    sc.add_triangles(soup)
    sc.commit()
    n = 1 << 16
    d_rays, d_hits = ctx.alloc(n * 32), ctx.alloc(n * 16)
    host = ctx.make_rays_uniform(0, n, device_buffer=d_rays)
    st = sc.trace_closest_device(d_rays, n, d_hits, ff.TRACE_BVH8 | ff.TRACE_COUNT)
    assert st["rays"] == n and st["nodes_visited"] > n and st["tris_tested"] > 0 and st["gpu_ms"] > 0
    got = sc.hits_unpack(d_hits, n)
    assert_hits_equal(got, sc.trace_closest(host, ff.TRACE_BVH8), "device entry vs host entry")
    sb = sc.trace_closest_device(d_rays, n, d_hits, ff.TRACE_BRUTE | ff.TRACE_COUNT)
    assert sb["tris_tested"] == n * 50000
    # bounce batch: cosine-hemisphere rays leaving the hits
    st = sc.trace_closest_device(d_rays, n, d_hits, ff.TRACE_BVH8)
    d_out = ctx.alloc(n * 32)
    nb = sc.make_rays_bounce(d_rays, d_hits, n, 77, d_out)
    assert nb == int((got["geom"] >= 0).sum())
    out = d_out.download(np.float32, nb * 8).reshape(2, nb, 4)
    assert np.allclose(np.linalg.norm(out[1, :, :3], axis=1), 1, atol=1e-5)
    assert (out[0, :, :3] > -0.1).all() and (out[0, :, :3] < 1.1).all()


# ----------------------------------------------------------------------------- binned-SAH builder (FF_BUILDER_SAH)
@pytest.mark.parametrize("n_tris", [1, 2, 3, 31, 32, 33, 34, 100, 1000, 8192, 8193, 12289, 20000, 200000])
def test_sah_builder_matches_brute_force(ff, ctx, n_tris):
    """The quality builder produces a different tree, never different hits: every size around the warp-kernel /
    CTA-kernel switch (32 | 33) and the one-CTA / many-CTA switch (8192 | 8193, chunk edges), BVH8 and BVH2 over the SAH tree against GPU brute force and against the LBVH tree."""
    soup = ctx.make_soup(n_tris, s=0.2 if n_tris < 1000 else None)
    sc = ctx.scene()
    sc.add_triangles(soup)
    sc.commit(ff.BUILDER_SAH)
    info = sc.info()
    assert info["n_triangles"] == n_tris and info["builder"] == ff.BUILDER_SAH and info["depth"] <= 32
    nr = 1 << 15
    rays = ctx.make_rays_uniform(0, nr, salt=9)
    brute = sc.trace_closest(rays, ff.TRACE_BRUTE)
    assert_hits_equal(sc.trace_closest(rays, ff.TRACE_BVH8), brute, "sah bvh8 vs brute")
    assert_hits_equal(sc.trace_closest(rays, ff.TRACE_BVH2), brute, "sah bvh2 vs brute")
    lb = ctx.scene()
    lb.add_triangles(soup)
    lb.commit(ff.BUILDER_LBVH)
    assert_hits_equal(lb.trace_closest(rays, ff.TRACE_BVH8), brute, "lbvh vs brute")
    if n_tris >= 20000:  # the point of the builder: a cheaper tree (fewer node visits per ray) on the soup
        n = 1 << 16
        d_rays, d_hits = ctx.alloc(n * 32), ctx.alloc(n * 16)
        ctx.make_rays_uniform(0, n, device_buffer=d_rays, host=False)
        a = sc.trace_closest_device(d_rays, n, d_hits, ff.TRACE_BVH8 | ff.TRACE_COUNT)
        b = lb.trace_closest_device(d_rays, n, d_hits, ff.TRACE_BVH8 | ff.TRACE_COUNT)
        assert a["nodes_visited"] < b["nodes_visited"], (a["nodes_visited"], b["nodes_visited"])
        assert info["sah_cost"] < lb.info()["sah_cost"]


def test_sah_builder_degenerate_inputs(ff, ctx):
    """All centres identical (no plane separates them: segments are halved by position) and exact duplicates."""
    tri = np.float32([[0, 0, 0, 1, 0, 0, 0, 1, 0]])
    same = np.repeat(tri, 300, 0)
    shifted = same.copy()
    shifted[150:, 2::3] = 0.5  # two stacks of coincident triangles
    rays = np.float32([[0.2, 0.2, 2, 0, 0, -1], [0.2, 0.2, 0.25, 0, 0, -1], [0.9, 0.9, 2, 0, 0, -1]])
    for tris in (same, shifted):
        sc = ctx.scene()
        sc.add_triangles(tris)
        sc.commit(ff.BUILDER_SAH)
        got = sc.trace_closest(rays, ff.TRACE_BVH8)
        assert_hits_equal(got, sc.trace_closest(rays, ff.TRACE_BRUTE), "degenerate sah")
        assert got["tri"][0] == (150 if tris is shifted else 0)  # lowest index among coincident hits (kernel.cu:115)


@pytest.mark.parametrize("cfg", ["C2", "C3"])
def test_sah_builder_config_scenes_vs_oracle(ff, ctx, orc, meshes, cfg):
    desc = S.config2(meshes, 64, 64) if cfg == "C2" else S.config3(meshes, 64, 36)
    sc, cam = S.instantiate(desc, ctx, meshes, builder=ff.BUILDER_SAH)
    osc = S.instantiate_on(desc, orc.scene(), meshes)
    rays = ctx.camera_rays(cam)
    assert_hits_equal(sc.trace_closest(rays), osc.intersect(rays, mode=1), "sah " + cfg)
    want, _ = osc.render(desc["camera"], 2, 3, seed=3)
    assert (bits(sc.render(cam, 2, 3, seed=3)) == bits(want)).all()


@pytest.mark.gpu
def test_clustered_scene_switches_to_63_bit_morton_keys(ff, ctx, monkeypatch):
    """A large scene whose triangles occupy a tiny part of its bounding box: on the 30-bit Morton grid nearly every triangle
    shares a cell with its neighbour, so the builder redoes codes + sort with 63-bit keys (ff_scene_info.morton_bits).  The
    hierarchy stays exact (BVH == brute force, bit for bit) and visits far fewer nodes than the one built from 30-bit keys."""
    rng = np.random.default_rng(63)
    n = (1 << 20) + 5000
    c = rng.random((n, 1, 3), dtype=np.float32) * np.float32(0.5)   # (triangles much smaller than this fall under the reference's det < 1e-6 rule)
    tris = (c + np.float32(5e-3) * (rng.random((n, 3, 3), dtype=np.float32) * 2 - 1)).reshape(n, 9).astype(np.float32)
    tris[-1] = np.array([900, 900, 900, 901, 900, 900, 900, 901, 900], np.float32)   # one far triangle stretches the bounds
    o = (rng.random((4096, 3), dtype=np.float32) * np.float32(0.5)).astype(np.float32)
    d = rng.normal(size=(4096, 3)).astype(np.float32)
    rays = np.concatenate([o, d], 1)
    sc = ctx.scene()
    sc.add_triangles(tris)
    sc.commit(ff.BUILDER_LBVH)
    assert sc.info()["morton_bits"] == 63
    got, st = sc.trace_closest(rays, mode=ff.TRACE_BVH8 | ff.TRACE_COUNT, want_stats=True)
    want = sc.trace_closest(rays, mode=ff.TRACE_BRUTE)
    assert (got["geom"] == want["geom"]).all() and (got["tri"] == want["tri"]).all() and (bits(got["t"]) == bits(want["t"])).all()
    assert got["hit"].sum() > 1000
    monkeypatch.setenv("FF_MORTON_BITS", "30")
    sc.commit(ff.BUILDER_LBVH)
    assert sc.info()["morton_bits"] == 30
    got30, st30 = sc.trace_closest(rays, mode=ff.TRACE_BVH8 | ff.TRACE_COUNT, want_stats=True)
    assert (got30["tri"] == want["tri"]).all() and (bits(got30["t"]) == bits(want["t"])).all()
    assert st["nodes_visited"] * 2 < st30["nodes_visited"], (st["nodes_visited"], st30["nodes_visited"])
    monkeypatch.delenv("FF_MORTON_BITS")
    # a uniform soup of the same size keeps the 30-bit keys: its duplicates are a few per cent
    soup = ctx.make_soup(n)
    sc2 = ctx.scene()
    sc2.add_triangles(soup)
    sc2.commit(ff.BUILDER_LBVH)
    assert sc2.info()["morton_bits"] == 30
    sc.close()
    sc2.close()


def test_host_batches_stream_through_one_launch(ff, ctx):
    """ff_trace_closest on host rays (the streaming launch: rays consumed while they are still being copied in, finished chunks
    copied out while it runs) == the device-resident entry point, bit for bit, for batches of several chunks with ragged ends,
    from page-locked and from pageable memory, on a scene with a hierarchy and on one without (rays finish in the fetch path)."""
    sc = ctx.scene()
    sc.add_triangles(ctx.make_soup(40000))
    sc.add_plane([0, 0, -0.5], [0, 0, 0], [4, 4, 4])
    sc.commit()
    empty = ctx.scene()
    empty.add_plane([0, 0, 0.5], [0, 0, 0], [1, 1, 1])
    empty.commit()
    n = (1 << 20) + (1 << 19) + 12345   # three full copy-in chunks (512 Ki rays) and a ragged fourth
    d_rays, d_hits = ctx.alloc(n * 32), ctx.alloc(n * 16)
    rays = ctx.make_rays_uniform(0, n, salt=5, device_buffer=d_rays)
    for scene in (sc, empty):
        scene.trace_closest_device(d_rays, n, d_hits)
        want = scene.hits_unpack(d_hits, n)
        assert 0 < want["hit"].sum() < n
        got = scene.trace_closest(rays)                       # pageable numpy memory
        assert_hits_equal(got, want, "host batch (pageable) vs device entry")
        h_rays = ctx.pinned_array((n, 6), np.float32, key="t_rays")
        h_rays[:] = rays
        g, t, k = ctx.pinned_array((n,), np.int32, key="t_g"), ctx.pinned_array((n,), np.int32, key="t_t"), ctx.pinned_array((n,), np.float32, key="t_k")
        for m in (n, (1 << 19) + 1, (1 << 19), 77):           # chunk boundary + 1, exactly one chunk, tiny
            g[:], t[:], k[:] = -9, -9, -9.0
            st = scene.trace_closest_into(h_rays[:m], g[:m], t[:m], k[:m])
            assert st["rays"] == m
            assert (g[:m] == want["geom"][:m]).all() and (t[:m] == want["tri"][:m]).all() and (bits(k[:m]) == bits(want["t"][:m])).all()
            assert (g[m:] == -9).all() and (k[m:] == -9.0).all()   # nothing written past the batch


def test_streaming_trace_falls_back_when_kernels_are_serialised_against_copies():
    """Under CUDA_LAUNCH_BLOCKING=1 (as under a profiler that replays kernels) the streaming launch can never see its rays: its
    lanes give up after a quarter of a second, and ff_trace_closest must redo the batch through the chunked pipeline and return the same
    hits -- twice, after which the context stops trying the streaming path.  Own process: the variable is read at CUDA start-up."""
    import os
    import subprocess
    import sys
    code = r'''
import sys, time
import numpy as np
sys.path.insert(0, %r)
import gpupathtracer_b200 as ff
if ff.device_count() == 0:
    print("NO_DEVICE"); sys.exit(0)
ctx = ff.Context(0)
sc = ctx.scene(); sc.add_triangles(ctx.make_soup(20000)); sc.commit()
n = 200000
d_rays, d_hits = ctx.alloc(n * 32), ctx.alloc(n * 16)
rays = ctx.make_rays_uniform(0, n, device_buffer=d_rays)
sc.trace_closest_device(d_rays, n, d_hits)
want = sc.hits_unpack(d_hits, n)
times = []
for _ in range(3):
    t0 = time.time()
    got = sc.trace_closest(rays)
    times.append(time.time() - t0)
    assert (got["geom"] == want["geom"]).all() and (got["tri"] == want["tri"]).all()
    assert (got["t"].view(np.uint32) == want["t"].view(np.uint32)).all()
print("OK", " ".join("%%.2f" %% t for t in times))
''' % (os.path.dirname(os.path.dirname(os.path.abspath(__file__))),)
    env = dict(os.environ, CUDA_LAUNCH_BLOCKING="1")
    r = subprocess.run([sys.executable, "-c", code], env=env, capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, r.stderr[-2000:]
    if "NO_DEVICE" in r.stdout:
        pytest.skip("no CUDA device")
    assert r.stdout.startswith("OK"), r.stdout
    t = [float(x) for x in r.stdout.split()[1:]]
    assert t[0] > 0.2 and t[1] > 0.2 and t[2] < 0.2, t   # two time-outs of 0.25 s, then the chunked pipeline straight away
