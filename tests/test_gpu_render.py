"""GPU: the integrators against the CPU oracle's as-intended path tracer (same Philox stream, same operation order,
samples summed in index order) -- the images are required to be BIT-IDENTICAL on one GPU, which is far inside the
per-pixel RMSE <= 1e-3 the north star asks for -- plus the as-coded normal frame, sample splitting and u8 output."""
import os

import numpy as np
import pytest

from conftest import bits
from gpupathtracer_b200 import multigpu as M
from gpupathtracer_b200 import scenes as S

pytestmark = pytest.mark.gpu


def descs(meshes):
    return {"C1": S.config1(96, 96), "C2": S.config2(meshes, 72, 72, 6, 4), "C3": S.config3(meshes, 96, 54, 5, 8)}


@pytest.mark.parametrize("cfg", ["C1", "C2", "C3"])
@pytest.mark.parametrize("aa", [0, 1])
def test_wavefront_and_megakernel_bit_identical_to_oracle(ff, ctx, orc, meshes, cfg, aa):
    desc = descs(meshes)[cfg]
    sc, cam = S.instantiate(desc, ctx, meshes)
    osc = S.instantiate_on(desc, orc.scene(), meshes)
    bg = (0.1, 0.2, 0.3) if cfg == "C1" else (0, 0, 0)
    want, want_rays = osc.render(desc["camera"], desc["spp"], desc["max_bounces"], seed=1234, aa=aa, background=bg)
    for integ in (ff.INTEGRATOR_WAVEFRONT, ff.INTEGRATOR_MEGAKERNEL):
        got, st = sc.render(cam, desc["spp"], desc["max_bounces"], want_stats=True, integrator=integ, seed=1234, aa=aa, background=bg)
        assert (bits(got) == bits(want)).all(), "%s integrator %d: max abs diff %g" % (cfg, integ, np.abs(got - want).max())
        assert st["rays"] == want_rays and st["paths"] == desc["spp"] * cam.width * cam.height
        rmse = float(np.sqrt(np.mean((got[..., :3] - want[..., :3]) ** 2)))
        assert rmse <= 1e-3  # the stated tolerance (trivially met: the images are identical)
    assert want[..., :3].max() > 0


@pytest.mark.parametrize("cfg", ["C2", "C3"])
def test_next_event_estimation_bit_identical_to_oracle(ff, ctx, orc, meshes, cfg):
    """The shadow-ray stage (wavefront: shade -> shadow queue -> any-hit kernel; megakernel: shadow phase of the lane
    state machine) against the oracle's NEE loop: same light samples, same visibility, sums in the same order."""
    desc = descs(meshes)[cfg]
    sc, cam = S.instantiate(desc, ctx, meshes)
    osc = S.instantiate_on(desc, orc.scene(), meshes)
    want, want_rays, want_shadow = osc.render(desc["camera"], desc["spp"], desc["max_bounces"], seed=77, aa=1, nee=1, want_shadow_rays=True)
    assert want_shadow > 0
    for integ in (ff.INTEGRATOR_WAVEFRONT, ff.INTEGRATOR_MEGAKERNEL):
        got, st = sc.render(cam, desc["spp"], desc["max_bounces"], want_stats=True, integrator=integ, seed=77, aa=1, nee=1)
        assert (bits(got) == bits(want)).all(), "%s integrator %d: max abs diff %g" % (cfg, integ, np.abs(got - want).max())
        assert st["rays"] == want_rays and st["shadow_rays"] == want_shadow
    off = sc.render(cam, desc["spp"], desc["max_bounces"], seed=77, aa=1, nee=0)
    assert (off != want).any()
    # passes of different size give the same image (the shadow queue is per pass)
    assert (bits(sc.render(cam, desc["spp"], desc["max_bounces"], seed=77, aa=1, nee=1, samples_per_pass=2)) == bits(want)).all()


def test_samples_per_pass_and_seed(ff, ctx, orc, meshes):
    desc = descs(meshes)["C2"]
    sc, cam = S.instantiate(desc, ctx, meshes)
    base = sc.render(cam, 6, 4, seed=99)
    for spb in (1, 2, 4, 6):
        assert (bits(sc.render(cam, 6, 4, seed=99, samples_per_pass=spb)) == bits(base)).all()
        assert (bits(sc.render(cam, 6, 4, seed=99, samples_per_pass=spb, integrator=ff.INTEGRATOR_MEGAKERNEL)) == bits(base)).all()
    other = sc.render(cam, 6, 4, seed=100)
    assert (other != base).any()
    big_seed = (0xABCDEF01 << 32) | 0x12345678
    osc = S.instantiate_on(desc, orc.scene(), meshes)
    want, _ = osc.render(desc["camera"], 2, 4, seed=big_seed)
    assert (bits(sc.render(cam, 2, 4, seed=big_seed)) == bits(want)).all()


def test_self_skip_of_the_origin_triangle_is_invisible(ff, ctx, orc, meshes):
    """Integrator rays never test the triangle they start on (make_ray_ctx, ff_trace.cuh); the CPU restatement tests every
    triangle.  Images must stay bit-identical for every ray offset: far above the guard of self_skip_safe (default 1e-4 in
    C2, whose global threshold pad/8 is 4.8e-6), just above it, below it and at zero (skip off), with and without next-event
    estimation, on flat (C2, C3), specular (showcase) and smooth-shaded (shaded) geometry."""
    all_descs = dict(descs(meshes), showcase=S.showcase(meshes, 64, 48, 2, 5), shaded=S.shaded(meshes, 64, 64, 2, 5))
    for name, spp, bounces in (("C2", 3, 4), ("C3", 2, 5), ("showcase", 2, 5), ("shaded", 2, 5)):
        desc = all_descs[name]
        sc, cam = S.instantiate(desc, ctx, meshes)
        osc = S.instantiate_on(desc, orc.scene(), meshes)
        for eps in (1e-4, 7e-5, 1e-5, 1e-6, 0.0):
            for nee in (0, 1):
                want, _ = osc.render(desc["camera"], spp, bounces, seed=5, ray_eps=eps, nee=nee)
                for integ in (ff.INTEGRATOR_WAVEFRONT, ff.INTEGRATOR_MEGAKERNEL):
                    got = sc.render(cam, spp, bounces, seed=5, ray_eps=eps, nee=nee, integrator=integ)
                    assert (bits(got) == bits(want)).all(), (name, eps, nee, integ)


def test_sample_split_matches_full_render(ff, ctx, meshes):
    """What the multi-GPU path does, on one GPU: disjoint sample ranges accumulated separately, reduced in rank order."""
    desc = descs(meshes)["C2"]
    sc, cam = S.instantiate(desc, ctx, meshes)
    spp, world = 8, 4
    full = sc.render(cam, spp, 4)
    npix = cam.width * cam.height
    bufs, partial = [], []
    for r in range(world):
        first, count = M.sample_range(r, world, spp)
        buf = ctx.alloc(npix * 16)
        buf.zero()
        sc.render_accumulate(cam, buf, count, 4, first_sample=first)
        bufs.append(buf)
        partial.append(buf.download(np.float32, npix * 4).reshape(cam.height, cam.width, 4))
    want = M.combine_sums(partial, spp)
    got = ctx.reduce_resolve_peers([b.ptr.value for b in bufs], cam.width, cam.height, spp)
    assert (bits(got) == bits(want)).all()  # fused reduce+resolve == host statement, bit for bit
    assert np.abs(got[..., :3] - full[..., :3]).max() <= 1e-5 * max(1.0, float(full.max()))  # vs in-order sum: fp32 re-association only
    # resuming: accumulate the second half into the first half's buffer == one render of all samples, bit for bit
    buf = ctx.alloc(npix * 16)
    buf.zero()
    sc.render_accumulate(cam, buf, 4, 4, first_sample=0)
    sc.render_accumulate(cam, buf, 4, 4, first_sample=4)
    assert (bits(ctx.resolve(buf, cam.width, cam.height, spp)) == bits(full)).all()


@pytest.mark.parametrize("cfg", ["C1", "C2", "C3", "main", "scaled"])
def test_as_coded_frame(ff, ctx, orc, meshes, golden, cfg):
    """ff_render_u8 / FF_INTEGRATOR_AS_CODED against the CPU restatement (bit for bit) and against the reference's own frame
    (oracle/_ref goldens): colour = |inverse(transpose(M)) * object-space unit normal| * 255 through make_uchar3's wrapping
    conversion (kernel.cu:117,183,214).  `scaled` has meshes under scales below and above 1."""
    desc = dict(descs(meshes), main=S.reference_main(160, 160), scaled=S.scaled_meshes(160, 160))[cfg]
    sc, cam = S.instantiate(desc, ctx, meshes)
    osc = S.instantiate_on(desc, orc.scene(), meshes)
    got = sc.render_u8(cam)
    assert (got == osc.render_as_coded(desc["camera"], mode=1)).all()
    as_float = sc.render(cam, 1, 0, integrator=ff.INTEGRATOR_AS_CODED)  # the same frame widened on the device
    assert (as_float[..., :3] == got.astype(np.float32) / np.float32(255)).all() and (as_float[..., 3] == 1).all()
    if cfg in ("main", "scaled"):  # against the reference's own frame: identical bytes wherever the same primitive is hit
        g = golden("ref_main.npz" if cfg == "main" else "ref_scaled.npz")
        hits = sc.trace_closest(g["rays"])
        same = (hits["geom"] == g["geom"]) & (hits["tri"] == g["tri"])
        assert same.mean() > 0.999  # baked vs object-space intersection may pick another triangle on a shared edge
        assert (got.reshape(-1, 3)[same] == g["as_coded"].reshape(-1, 3)[same]).all()


def test_render_u8_and_device_output(ff, ctx, meshes):
    desc = descs(meshes)["C2"]
    sc, cam = S.instantiate(desc, ctx, meshes)
    img = sc.render(cam, 4, 4)
    u8 = sc.render_u8(cam, 4, 4, integrator=ff.INTEGRATOR_WAVEFRONT)
    want = (np.clip(img[..., :3], 0, 1) * np.float32(255)).astype(np.uint8)
    assert (u8 == want).all()
    d_out = ctx.alloc(cam.width * cam.height * 16)
    sc.render(cam, 4, 4, host=False, device_out=d_out)
    assert (bits(d_out.download(np.float32, cam.width * cam.height * 4)) == bits(img).ravel()).all()


@pytest.mark.parametrize("integ", ["wavefront", "megakernel"])
@pytest.mark.parametrize("world", [2, 3, 8])
def test_tile_split_sums_to_the_full_frame_bit_exactly(ff, ctx, meshes, integ, world):
    """Interleaved-tile split (SURVEY 8e ii): rank k renders only its tiles' pixels; the SUM of the ranks' accumulators is
    bit-identical to the single-GPU accumulator, for a frame that is not a multiple of the tile size, both integrators,
    NEE on (the shadow queue indexes paths, not pixels)."""
    integrator = ff.INTEGRATOR_WAVEFRONT if integ == "wavefront" else ff.INTEGRATOR_MEGAKERNEL
    desc = S.config2(meshes, 75, 52, 5, 4)
    sc, cam = S.instantiate(desc, ctx, meshes)
    npix = cam.width * cam.height
    kw = dict(integrator=integrator, seed=9, aa=1, nee=1)
    whole = ctx.alloc(npix * 16)
    whole.zero()
    st_full = sc.render_accumulate(cam, whole, 5, 4, **kw)
    want = whole.download(np.float32, npix * 4)
    total = np.zeros(npix * 4, np.float32)
    paths = rays = 0
    for r in range(world):
        part = ctx.alloc(npix * 16)
        part.zero()
        st = sc.render_accumulate(cam, part, 5, 4, tile_rank=r, tile_world=world, tile_size=16, **kw)
        got = part.download(np.float32, npix * 4).reshape(npix, 4)
        own = ff.tile_owned_pixels(cam.width, cam.height, r, world, 16)
        mask = np.ones(npix, bool)
        mask[own] = False
        assert (got[mask] == 0).all()   # pixels of other ranks are untouched
        assert st["paths"] == 5 * len(own)
        total = total + got.ravel()
        paths += st["paths"]
        rays += st["rays"]
        part.free()
    assert (bits(total) == bits(want)).all()
    assert paths == st_full["paths"] and rays == st_full["rays"]
    whole.free()


@pytest.mark.parametrize("integ", ["wavefront", "megakernel"])
def test_progressive_accumulation_with_camera_moves(ff, ctx, meshes, integ):
    """ff_progressive (the viewer loop's frame-to-frame state, kernel.cu:331-359): k frames of s samples are bit-identical
    to one render of k*s samples; a camera control (Camera::ProcessKeyboard, utilities.h:343-382) restarts the
    accumulation at the new pose; an unchanged camera keeps accumulating; the 8-bit frame follows the PBO contract."""
    integrator = ff.INTEGRATOR_WAVEFRONT if integ == "wavefront" else ff.INTEGRATOR_MEGAKERNEL
    desc = descs(meshes)["C2"]
    sc, cam = S.instantiate(desc, ctx, meshes)
    prog = ff.Progressive(sc)
    kw = dict(integrator=integrator, seed=5, aa=1)
    for k in range(1, 4):
        img, total, u8 = prog.frame(cam, 2, 4, want_u8=True, **kw)
        assert total == 2 * k
        want = sc.render(cam, 2 * k, 4, **kw)
        assert (bits(img) == bits(want)).all(), "frame %d" % k
        assert (u8 == (np.clip(want[..., :3], 0, 1) * np.float32(255)).astype(np.uint8)).all()
    # move: the accumulation restarts at the new pose
    cam.process_keyboard(ff.CAM_FORWARD).process_keyboard(ff.CAM_YAW_LEFT).process_mouse(12.0, -7.0)
    img, total = prog.frame(cam, 3, 4, **kw)
    assert total == 3 and (bits(img) == bits(sc.render(cam, 3, 4, **kw))).all()
    img, total = prog.frame(cam, 1, 4, **kw)
    assert total == 4 and (bits(img) == bits(sc.render(cam, 4, 4, **kw))).all()
    # an image-changing parameter restarts it too; an explicit reset as well
    img, total = prog.frame(cam, 2, 3, **kw)
    assert total == 2 and (bits(img) == bits(sc.render(cam, 2, 3, **kw))).all()
    prog.reset()
    img, total = prog.frame(cam, 2, 3, **kw)
    assert total == 2 and (bits(img) == bits(sc.render(cam, 2, 3, **kw))).all()
    # a different image size reallocates
    small = ff.Camera.from_dict(dict(desc["camera"], width=40, height=24))
    img, total = prog.frame(small, 2, 3, **kw)
    assert img.shape == (24, 40, 4) and total == 2 and (bits(img) == bits(sc.render(small, 2, 3, **kw))).all()
    prog.close()


def test_errors(ff, ctx, meshes):
    sc = ctx.scene()
    sc.add_triangles(meshes["cube"])
    cam = ff.Camera(width=8, height=8)
    with pytest.raises(ff.FireflyError):
        sc.render(cam, 1, 1)  # not committed
    sc.commit()
    with pytest.raises(ff.FireflyError):
        sc.render(cam, 0, 1)
    with pytest.raises(ff.FireflyError):
        sc.render(cam, 1, 1, nee=2)
    assert sc.render(cam, 1, 1, nee=1).shape == (8, 8, 4)  # no emitter in the scene: NEE has nothing to sample
    with pytest.raises(ff.FireflyError):
        sc.trace_closest(np.zeros((4, 6), np.float32), mode=77)
    with pytest.raises(ff.FireflyError):
        ff.Context(10 ** 6)


def test_scene_outlives_context(ff, meshes):
    """Destroying a context releases its scenes' device memory; the orphaned handle fails cleanly and frees safely."""
    if ff.device_count() == 0:
        pytest.skip("no CUDA device")
    c = ff.Context(0)
    sc = c.scene()
    sc.add_triangles(meshes["cube"])
    sc.commit()
    c.close()
    with pytest.raises(ff.FireflyError):
        sc.trace_closest(np.zeros((1, 6), np.float32))
    with pytest.raises(ff.FireflyError):
        sc.commit()
    sc.close()


@pytest.mark.parametrize("nee", [0, 1])
def test_showcase_all_material_and_geometry_types(ff, ctx, orc, meshes, nee):
    """Every BXDFTyp (emitter, diffuse, mirror, glass) and GeometryType (sphere, plane, mesh) in one frame: both
    integrators bit-identical to the oracle, closest-hit ids and any-hit answers included."""
    desc = S.showcase(meshes, 80, 64, 4, 6)
    sc, cam = S.instantiate(desc, ctx, meshes)
    osc = S.instantiate_on(desc, orc.scene(), meshes)
    rays = ctx.camera_rays(cam)
    got, want = sc.trace_closest(rays), osc.intersect(rays, mode=1)
    assert (got["geom"] == want["geom"]).all() and (got["tri"] == want["tri"]).all() and (bits(got["t"]) == bits(want["t"])).all()
    assert set(np.unique(want["geom"])) >= {0, 2, 3, 5}  # walls, glass mesh, mirror sphere, plane are all visible
    assert (sc.trace_closest(rays, ff.TRACE_BRUTE)["geom"] == want["geom"]).all()
    assert (sc.trace_any(rays, None) == (want["geom"] >= 0)).all()
    img, rays_o, shadow_o = osc.render(desc["camera"], desc["spp"], desc["max_bounces"], seed=11, aa=1, nee=nee, want_shadow_rays=True)
    for integ in (ff.INTEGRATOR_WAVEFRONT, ff.INTEGRATOR_MEGAKERNEL):
        out, st = sc.render(cam, desc["spp"], desc["max_bounces"], want_stats=True, integrator=integ, seed=11, aa=1, nee=nee)
        assert (bits(out) == bits(img)).all(), "integrator %d nee %d: max abs diff %g" % (integ, nee, np.abs(out - img).max())
        assert st["rays"] == rays_o and st["shadow_rays"] == shadow_o
    assert img[..., :3].max() > 0


def test_scene_file_and_headless_cli(ff, ctx, orc, meshes, tmp_path):
    """Scene file -> library (ff_scene_load_file) and -> firefly_cli (C++ over the C ABI only) -> PFM: both give the frame
    the Python-built description gives, which is the oracle's frame."""
    import json
    import subprocess
    desc = S.showcase(meshes, 64, 48, 3, 5)
    path = S.write_scene_file(desc, str(tmp_path / "show.ffscene"), meshes, render=dict(seed=21, aa=1, nee=1))
    osc = S.instantiate_on(desc, orc.scene(), meshes)
    want, _ = osc.render(desc["camera"], 3, 5, seed=21, aa=1, nee=1)
    sc, cam, d = ff.load_scene_file(ctx, path)
    sc.commit(d.builder)
    got = sc.render(cam, d.params.spp, d.params.max_bounces, seed=d.params.seed, aa=d.params.aa, nee=d.params.nee)
    assert (bits(got) == bits(want)).all()
    cli = os.path.join(os.path.dirname(ff.__file__), "firefly_cli")
    out = str(tmp_path / "show.pfm")
    r = subprocess.run([cli, path, "--out", out, "--integrator", "megakernel"], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr
    st = json.loads(r.stdout.strip().splitlines()[-1])
    assert st["geometries"] == 6 and st["spp"] == 3 and st["nee"] == 1 and st["rays"] > 0 and st["shadow_rays"] > 0
    raw = open(out, "rb").read()
    hdr, dims, scale, body = raw.split(b"\n", 3)
    assert hdr == b"PF" and dims.split() == [b"64", b"48"] and float(scale) < 0
    img = np.frombuffer(body, "<f4").reshape(48, 64, 3)[::-1]  # PFM rows run bottom to top
    assert (bits(img) == bits(want[..., :3])).all()
    png = str(tmp_path / "show.png")
    r = subprocess.run([cli, path, "--out", png, "--spp", "1", "--size", "32", "24", "--builder", "sah"], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and open(png, "rb").read(8) == b"\x89PNG\r\n\x1a\n"
    r = subprocess.run([cli, str(tmp_path / "nope.ffscene")], capture_output=True, text=True, timeout=60)
    assert r.returncode == 1 and "cannot read scene file" in r.stderr


@pytest.fixture(scope="module")
def attrs():
    return S.load_mesh_attrs()


@pytest.mark.gpu
@pytest.mark.parametrize("nee", [0, 1])
def test_shading_attributes_bit_identical_to_oracle(ff, ctx, orc, meshes, attrs, nee):
    """interpolated corner normals (diffuse, mirror, glass; one mesh under a non-uniform scale), uvs and
    per-face materials (emissive faces join the NEE light list): both integrators bit-identical to the oracle, ray and
    shadow-ray counts included; every feature buffer bit-identical; hit ids unaffected by the attributes."""
    desc = S.shaded(meshes, 72, 56, 4, 5)
    sc, cam = S.instantiate(desc, ctx, meshes, attrs=attrs)
    osc = S.instantiate_on(desc, orc.scene(), meshes, attrs)
    rays = ctx.camera_rays(cam)
    got, want = sc.trace_closest_uv(rays), osc.intersect_uv(rays)
    assert (got["geom"] == want["geom"]).all() and (got["tri"] == want["tri"]).all()
    assert (bits(got["u"]) == bits(want["u"])).all() and (bits(got["v"]) == bits(want["v"])).all()
    assert set(np.unique(want["geom"])) >= {0, 2, 3, 4, 5, 6}
    if nee == 0:
        for kind in (ff.AOV_NORMAL, ff.AOV_UV, ff.AOV_BARY, ff.AOV_ALBEDO):
            a, b = sc.render(cam, integrator=kind), osc.render_aov(desc["camera"], kind)
            assert (bits(a) == bits(b)).all(), "feature buffer %d: max abs diff %g" % (kind, np.abs(a - b).max())
        assert len(np.unique(sc.render(cam, integrator=ff.AOV_ALBEDO).reshape(-1, 4), axis=0)) >= 5  # per-face materials visible
    img, rays_o, shadow_o = osc.render(desc["camera"], desc["spp"], desc["max_bounces"], seed=17, aa=1, nee=nee, want_shadow_rays=True)
    for integ in (ff.INTEGRATOR_WAVEFRONT, ff.INTEGRATOR_MEGAKERNEL):
        out, st = sc.render(cam, desc["spp"], desc["max_bounces"], want_stats=True, integrator=integ, seed=17, aa=1, nee=nee)
        assert (bits(out) == bits(img)).all(), "integrator %d nee %d: max abs diff %g" % (integ, nee, np.abs(out - img).max())
        assert st["rays"] == rays_o and st["shadow_rays"] == shadow_o
    assert img[..., :3].max() > 0
    # the SAH tree shades identically (attributes are addressed by flat index, not by leaf slot)
    sc.commit(ff.BUILDER_SAH)
    out = sc.render(cam, desc["spp"], desc["max_bounces"], seed=17, aa=1, nee=nee)
    assert (bits(out) == bits(img)).all()


@pytest.mark.gpu
def test_obj_attributes_through_the_scene(ff, ctx, orc, meshes, attrs, tmp_path):
    """ff_scene_add_obj_ex: the attributes of an OBJ + MTL pair reach shading (flags pick which), a missing library or an
    unknown material name keeps the geometry's material, flip swaps normals and uvs with the corners; setter errors."""
    from test_host import _write_obj_full
    name = "sphere"
    xyz9, n9 = meshes[name], attrs[name + "_n"]
    n = len(xyz9)
    uv6 = np.stack([xyz9[:, 0], xyz9[:, 1], xyz9[:, 3], xyz9[:, 4], xyz9[:, 6], xyz9[:, 7]], 1).astype(np.float32) * np.float32(0.25)
    face_mat = ((np.arange(n) // 40) % 3).astype(np.int32)
    (tmp_path / "lib.mtl").write_text("newmtl red\nKd 0.9 0.1 0.1\n\nnewmtl mirror\nKs 0.8 0.8 0.9\nillum 3\n\nnewmtl lamp\nKe 2 1.5 1\n")
    mats = [dict(type=1, albedo=(0.9, 0.1, 0.1), specular=(1, 1, 1)), dict(type=2, albedo=(0.8, 0.8, 0.8), specular=(0.8, 0.8, 0.9)),
            dict(type=0, albedo=(0.8, 0.8, 0.8), emissive=(2, 1.5, 1), intensity=1.0)]
    p = str(tmp_path / "ball.obj")
    _write_obj_full(p, xyz9, n9, uv6, face_mat, ["red", "mirror", "unknown"], mtl="lib.mtl")
    face_mat_known = np.where(face_mat == 2, -1, face_mat).astype(np.int32)   # "unknown" is not in the library
    own = dict(type=1, albedo=(0.3, 0.6, 0.9))
    cam_d = S.camera((0, 0.5, 9), fov=40.0, width=48, height=40)
    cam = ff.Camera.from_dict(cam_d)
    box = S.flip_winding(meshes["cube"])

    def build(flags, flip=False):
        sc = ctx.scene()
        sc.add_triangles(box, scale=(12, 12, 24), material=S.EMITTER_WHITE2)
        sc.add_obj_ex(p, material=own, flags=flags | (ff.OBJ_FLIP_WINDING if flip else 0))
        sc.commit()
        return sc

    def oracle(normals, uvs, fm, flip=False):
        sc = orc.scene()
        sc.add_mesh(box, scale=(12, 12, 24), material=S.EMITTER_WHITE2)
        g = S.geom(xyz9, material=own, flip=flip, normals=n9 if normals else None, uvs=uv6 if uvs else None,
                   face_materials=(face_mat_known, mats[:2]) if fm else None)
        gi = sc.add_mesh(S.resolve_mesh(g, None), material=own)
        S._apply_attrs(sc, gi, g, None)
        return sc

    for flags, flip in ((0, False), (ff.OBJ_NORMALS, False), (ff.OBJ_NORMALS | ff.OBJ_UVS | ff.OBJ_MATERIALS, False),
                        (ff.OBJ_NORMALS | ff.OBJ_UVS | ff.OBJ_MATERIALS, True)):
        sc = build(flags, flip)
        osc = oracle(flags & ff.OBJ_NORMALS, flags & ff.OBJ_UVS, flags & ff.OBJ_MATERIALS, flip)
        for kind in (ff.AOV_NORMAL, ff.AOV_UV, ff.AOV_ALBEDO):
            assert (bits(sc.render(cam, integrator=kind)) == bits(osc.render_aov(cam_d, kind))).all(), (flags, flip, kind)
        want, _ = osc.render(cam_d, 2, 3, seed=4, aa=1)
        assert (bits(sc.render(cam, 2, 3, seed=4, aa=1)) == bits(want)).all(), (flags, flip)
    # the same through a scene file: `mesh ... normals uvs materials`
    sf = tmp_path / "ball.ffscene"
    sf.write_text("camera pos 0 0.5 9 yaw -90 fov 40 size 48 40 near 0.1 far 1000\n"
                  "material own diffuse albedo 0.3 0.6 0.9\nmaterial lamp emitter emissive 1 1 1 intensity 2\n"
                  "triangles scale 12 12 24 material lamp flip\n" + "".join("tri %s\n" % " ".join("%.9g" % v for v in t) for t in meshes["cube"]) +
                  "end\nmesh ball.obj material own normals uvs materials\nrender spp 2 bounces 3 seed 4 aa 1\n")
    fsc, fcam, d = ff.load_scene_file(ctx, str(sf))
    fsc.commit(d.builder)
    osc = oracle(True, True, True)
    for kind in (ff.AOV_NORMAL, ff.AOV_UV, ff.AOV_ALBEDO):
        assert (bits(fsc.render(fcam, integrator=kind)) == bits(osc.render_aov(cam_d, kind))).all(), kind
    # a missing material library is an I/O error only when materials were asked for
    q = str(tmp_path / "nolib.obj")
    _write_obj_full(q, xyz9, n9, uv6, face_mat, ["red", "mirror", "lamp"], mtl="absent.mtl")
    sc = ctx.scene()
    sc.add_obj_ex(q, flags=ff.OBJ_NORMALS)
    with pytest.raises(ff.FireflyError):
        sc.add_obj_ex(q, flags=ff.OBJ_MATERIALS)
    with pytest.raises(ff.FireflyError):
        sc.add_obj_ex(q, flags=64)
    gi = sc.add_plane()
    with pytest.raises(ff.FireflyError):
        sc.set_normals(gi, n9)
    with pytest.raises(ff.FireflyError):
        sc.set_face_materials(0, np.full(n, 3, np.int32), mats)   # index 3 of 3 materials
    with pytest.raises(ff.FireflyError):
        sc.set_uvs(7, uv6)


@pytest.mark.gpu
def test_converged_renders_agree_within_monte_carlo_tolerance(ff, ctx, meshes):
    """North-star correctness part 3 on the GPU at a converged sample count: config 2 at 128 x 128, 4096 spp.  The plain loop
    and next-event estimation are two unbiased estimators of the same image, so their means must agree within the noise that
    two independent seeds of each estimator show.  Tolerances (written here, as the contract asks): per 8 x 8-pixel block the
    difference of the two estimators' means is measured in units of its own standard error, estimated from the seed pairs --
    |z| of the WHOLE-IMAGE mean below 4, fewer than 2 % of the blocks beyond 6 standard errors (seed-pair variances have one
    degree of freedom, hence the wide per-block bar), image means within 1 %.  Sample-split accumulation at this size stays
    within the fp32 re-association bound of 1e-5 relative."""
    w = h = 128
    spp, bounces = 4096, 4
    desc = S.config2(meshes, w, h, spp, bounces)
    sc, cam = S.instantiate(desc, ctx, meshes, builder=1)
    img = {}
    for nee in (0, 1):
        for seed in (11, 2222):
            img[nee, seed] = sc.render(cam, spp, bounces, nee=nee, seed=seed, aa=1)[..., :3].astype(np.float64)

    def blocks(a):
        return a.reshape(h // 8, 8, w // 8, 8, 3).mean(axis=(1, 3, 4))
    m0, m1 = blocks((img[0, 11] + img[0, 2222]) / 2), blocks((img[1, 11] + img[1, 2222]) / 2)
    v0, v1 = blocks(img[0, 11] - img[0, 2222]) ** 2 / 4, blocks(img[1, 11] - img[1, 2222]) ** 2 / 4   # variance of a two-seed mean
    # block variances from ONE seed pair are noisy: pool them over the image for the per-block bar, keep them for the global one
    se_block = np.sqrt(np.mean(v0 + v1))
    z_blocks = (m1 - m0) / se_block
    z_global = (m1 - m0).sum() / np.sqrt((v0 + v1).sum())
    assert m0.mean() > 0.05, "the frame is not lit"
    assert abs(z_global) < 4.0, z_global
    assert (np.abs(z_blocks) > 6.0).mean() < 0.02, np.abs(z_blocks).max()
    assert abs(m1.mean() - m0.mean()) < 0.01 * m0.mean(), (m0.mean(), m1.mean())
    # disjoint sample ranges summed == one render of all samples (Philox is keyed by the global sample index)
    npix = w * h
    buf = ctx.alloc(npix * 16)
    buf.zero()
    for k in range(4):
        sc.render_accumulate(cam, buf, spp // 4, bounces, first_sample=k * (spp // 4), seed=11, aa=1)
    split = ctx.resolve(buf, w, h, spp)[..., :3]
    one = img[0, 11]
    assert np.abs(split - one).max() <= 1e-5 * max(1.0, float(one.max()))
