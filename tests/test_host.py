"""CPU: the host side of the C ABI -- library loads and exports what include/firefly.h declares, OBJ reader, camera and
geometry matrices (bit-exact against the oracle), image writers, error behaviour.  No kernel is launched here."""
import ctypes
import os
import re
import struct
import zlib

import numpy as np
import pytest

from conftest import ROOT, bits
from gpupathtracer_b200 import scenes as S


def test_library_exports_every_declared_symbol(ff):
    hdr = open(os.path.join(ROOT, "include", "firefly.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    names = sorted(set(re.findall(r"\b(ff_[a-z0-9_]+)\s*\(", hdr)))
    assert len(names) >= 40
    lib = ctypes.CDLL(ff.LIB_PATH)
    missing = [n for n in names if not hasattr(lib, n)]
    assert not missing, missing
    assert lib.ff_version() == 100


def test_no_cpu_fallback(ff):
    if ff.device_count() > 0:
        pytest.skip("a CUDA device is present")
    with pytest.raises(ff.FireflyError) as e:
        ff.Context(0)
    assert "no CPU fallback" in str(e.value)


def _write_obj(path, tris, relative=False, with_extras=False, quads=None):
    v = tris.reshape(-1, 3)
    with open(path, "w") as f:
        f.write("# written by tests\nmtllib missing.mtl\no thing\n")
        for p in v:
            f.write("v %r %r %r\n" % (float(p[0]), float(p[1]), float(p[2])))
        if with_extras:
            f.write("vn 0 0 1\nvt 0.5 0.5\ns off\nusemtl none\n")
        n = len(v)
        for i in range(len(tris)):
            idx = [3 * i + 1, 3 * i + 2, 3 * i + 3]
            if relative:
                idx = [k - n - 1 for k in idx]
            if with_extras:
                f.write("f %d/1/1 %d/1/1 %d//1\n" % tuple(idx))
            else:
                f.write("f %d %d %d\n" % tuple(idx))
        if quads is not None:
            base = n
            for q in quads:
                for p in q:
                    f.write("v %r %r %r\n" % tuple(float(x) for x in p))
                f.write("f %d %d %d %d\n" % (base + 1, base + 2, base + 3, base + 4))
                base += 4


def test_obj_reader(ff, meshes, tmp_path):
    for name in ("cube", "triangle", "sphere"):
        for relative in (False, True):
            for extras in (False, True):
                p = str(tmp_path / ("%s_%d_%d.obj" % (name, relative, extras)))
                _write_obj(p, meshes[name], relative, extras)
                got = ff.load_obj(p)
                assert got.shape == meshes[name].shape
                assert (bits(got) == bits(meshes[name])).all()
    quad = np.array([[[0, 0, 0], [1, 0, 0], [1, 1, 0], [0, 1, 0]]], np.float32)
    p = str(tmp_path / "quad.obj")
    _write_obj(p, meshes["triangle"], quads=quad)
    got = ff.load_obj(p)
    assert got.shape == (3, 9)  # 1 triangle + fan of the quad
    assert (got[1] == [0, 0, 0, 1, 0, 0, 1, 1, 0]).all() and (got[2] == [0, 0, 0, 1, 1, 0, 0, 1, 0]).all()
    with pytest.raises(ff.FireflyError):
        ff.load_obj(str(tmp_path / "does_not_exist.obj"))
    bad = tmp_path / "bad.obj"
    bad.write_text("v 0 0 0\nv 1 0 0\nf 1 2 7\n")
    with pytest.raises(ff.FireflyError):
        ff.load_obj(str(bad))
    empty = tmp_path / "empty.obj"
    empty.write_text("# nothing\n")
    assert ff.load_obj(str(empty)).shape == (0, 9)


@pytest.mark.skipif(not os.path.isdir("/root/reference/sceneResources"), reason="reference tree not mounted")
def test_obj_reader_on_reference_assets(ff, meshes):
    for name in S.MESH_NAMES:
        got = ff.load_obj("/root/reference/sceneResources/%s.obj" % name)
        assert (bits(got) == bits(meshes[name])).all(), name


def test_camera_matches_oracle_bit_exact(ff, orc, golden):
    g = golden("ref_camera.npz")
    for i in range(3):
        c = g["cam%d" % i]
        cam = ff.Camera(c[0:3], c[3], c[4], c[5], int(c[6]), int(c[7]), c[8], c[9])
        f, r, u = cam.basis()
        assert (bits(np.stack([f, r, u])) == bits(g["basis%d" % i])).all()
        iv, ip = cam.inverse_matrices()
        assert (bits(iv) == bits(g["iv%d" % i])).all() and (bits(ip) == bits(g["ip%d" % i])).all()
    rng = np.random.default_rng(5)
    for _ in range(50):
        d = S.camera(rng.uniform(-5, 5, 3), yaw=rng.uniform(-180, 180), pitch=rng.uniform(-80, 80), fov=rng.uniform(20, 100),
                     width=int(rng.integers(1, 500)), height=int(rng.integers(1, 500)))
        iv, ip = ff.Camera.from_dict(d).inverse_matrices()
        oiv, oip = orc.camera_inverse_matrices(d)
        assert (bits(iv) == bits(oiv)).all() and (bits(ip) == bits(oip)).all()


def test_philox_host(ff, orc):
    rng = np.random.default_rng(11)
    for _ in range(100):
        ctr = rng.integers(0, 2 ** 32, 4, dtype=np.uint64).astype(np.uint32)
        key = rng.integers(0, 2 ** 32, 2, dtype=np.uint64).astype(np.uint32)
        assert (ff.philox4x32_10(ctr, key) == orc.philox(ctr, key)).all()
    assert [int(x) for x in ff.philox4x32_10([0] * 4, [0] * 2)] == [0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8]


def _png_decode(path):
    b = open(path, "rb").read()
    assert b[:8] == b"\x89PNG\r\n\x1a\n"
    pos, chunks = 8, {}
    while pos < len(b):
        n, typ = struct.unpack(">I4s", b[pos:pos + 8])
        data = b[pos + 8:pos + 8 + n]
        crc, = struct.unpack(">I", b[pos + 8 + n:pos + 12 + n])
        assert crc == zlib.crc32(typ + data) & 0xffffffff
        chunks.setdefault(typ, b"")
        chunks[typ] += data
        pos += 12 + n
    w, h, depth, ctype = struct.unpack(">IIBB", chunks[b"IHDR"][:10])
    assert depth == 8 and ctype == 2
    raw = np.frombuffer(zlib.decompress(chunks[b"IDAT"]), np.uint8).reshape(h, 1 + 3 * w)
    assert (raw[:, 0] == 0).all()
    return raw[:, 1:].reshape(h, w, 3)


def test_image_writers(ff, tmp_path):
    rng = np.random.default_rng(2)
    img = rng.uniform(0, 1.2, (37, 53, 4)).astype(np.float32)
    u8 = rng.integers(0, 256, (37, 53, 3), dtype=np.uint8)
    ff.write_image(str(tmp_path / "a.png"), u8)
    assert (_png_decode(str(tmp_path / "a.png")) == u8).all()
    ff.write_image(str(tmp_path / "a.ppm"), u8)
    b = open(tmp_path / "a.ppm", "rb").read()
    assert b.startswith(b"P6\n53 37\n255\n") and np.frombuffer(b[len(b) - u8.size:], np.uint8).tobytes() == u8.tobytes()
    ff.write_image(str(tmp_path / "f.png"), img)
    dec = _png_decode(str(tmp_path / "f.png")).astype(np.float64) / 255
    lin = np.clip(img[..., :3], 0, 1).astype(np.float64)
    srgb = np.where(lin <= 0.0031308, 12.92 * lin, 1.055 * lin ** (1 / 2.4) - 0.055)
    assert np.abs(dec - srgb).max() <= 0.5 / 255 + 1e-6
    ff.write_image(str(tmp_path / "f.pfm"), img)
    b = open(tmp_path / "f.pfm", "rb").read()
    hdr = b"PF\n53 37\n-1.0\n"
    assert b.startswith(hdr)
    pf = np.frombuffer(b[len(hdr):], np.float32).reshape(37, 53, 3)[::-1]
    assert (pf == img[..., :3]).all()
    ff.write_image(str(tmp_path / "f.exr"), img)
    b = open(tmp_path / "f.exr", "rb").read()
    assert struct.unpack("<I", b[:4])[0] == 20000630 and b"channels\x00chlist\x00" in b
    # last scanline block: y, byte count, then B, G, R planes
    line = 53 * 3 * 4
    blk = b[len(b) - (8 + line):]
    y, nbytes = struct.unpack("<ii", blk[:8])
    assert y == 36 and nbytes == line
    planes = np.frombuffer(blk[8:], np.float32).reshape(3, 53)
    assert (planes[2] == img[36, :, 0]).all() and (planes[0] == img[36, :, 2]).all()
    with pytest.raises(ff.FireflyError):
        ff.write_image(str(tmp_path / "f.bmp"), img)


def test_scene_descriptions(meshes):
    top, rest = S._cube_faces(meshes["cube"])
    assert len(top) == 2 and len(rest) == 10
    d2 = S.config2(meshes)
    assert d2["camera"]["width"] == 1024 and d2["spp"] == 64 and d2["max_bounces"] == 4
    assert sum(len(S.resolve_mesh(g, meshes)) for g in d2["geoms"]) == 6048
    d3 = S.config3(meshes)
    assert sum(len(S.resolve_mesh(g, meshes)) for g in d3["geoms"]) == 5172 + 320 + 2
    f = S.flip_winding(meshes["triangle"])
    assert (f[0] == [0, 0, 0, 0, 1, 0, 1, 0, 0]).all()


def test_scene_file_parser(ff, meshes, tmp_path):
    """The scene description format (replaces the literals of main(), kernel.cu:227-259): parse without a device."""
    p = S.write_scene_file(S.showcase(meshes, 96, 64, 3, 5), str(tmp_path / "show.ffscene"), meshes,
                           render=dict(seed=7, nee=1, aa=1, integrator="megakernel", builder="sah", background=(0.1, 0.2, 0.3)))
    d = ff.validate_scene_file(p)
    assert d.n_geometries == 6 and d.builder == ff.BUILDER_SAH
    assert (d.camera.width, d.camera.height, d.camera.fov_deg, d.camera.yaw_deg) == (96, 64, 70.0, -90.0)
    assert (d.params.spp, d.params.max_bounces, d.params.seed, d.params.nee, d.params.aa) == (3, 5, 7, 1, 1)
    assert d.params.integrator == ff.INTEGRATOR_MEGAKERNEL and np.allclose(list(d.params.background), [0.1, 0.2, 0.3])
    assert os.path.exists(tmp_path / "rocketman.obj")
    assert (ff.load_obj(str(tmp_path / "rocketman.obj")).view(np.uint32) == meshes["rocketman"].view(np.uint32)).all()
    bad = {"camera pos 1 2": "pos needs 3", "mesh": "needs a path", "sphere -1": "positive radius", "frobnicate 1": "unknown directive",
           "plane material nope": "unknown material", "triangles\ntri 1 2 3": "exactly 9", "triangles\ntri 0 0 0 1 0 0 0 1 0": "without 'end'",
           "render spp 0": "positive integer", "camera size 10 10": "no geometry", "material a shiny": "unknown material type"}
    for text, msg in bad.items():
        q = tmp_path / "bad.ffscene"
        q.write_text(text + "\n")
        with pytest.raises(ff.FireflyError) as e:
            ff.validate_scene_file(str(q))
        assert msg in str(e.value), (text, str(e.value))
    with pytest.raises(ff.FireflyError):
        ff.validate_scene_file(str(tmp_path / "missing.ffscene"))
