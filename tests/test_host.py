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
def test_obj_reader_short_lines_never_borrow_from_the_next_line(ff, tmp_path):
    """A number parse is bounded by its own line: a one-coordinate `vt` is legal (v = 0), a short `v` line is an error even
    when the next line starts with a number, and a file whose last number ends exactly on a page boundary loads."""
    p = tmp_path / "short_vt.obj"
    p.write_text("v 0 0 0\nv 1 0 0\nv 0 1 0\nvt 0.25\n0.5 0.5\nvt 0.5 0.75\nvt 1 1\nf 1/1 2/2 3/3\n")
    m = ff.load_obj_ex(str(p), threads=1)
    assert m["xyz9"].shape == (1, 9) and m["uvs6"].tolist() == [[0.25, 0.0, 0.5, 0.75, 1.0, 1.0]]
    bad = tmp_path / "short_v.obj"
    bad.write_text("v 0 0 0\nv 1 2\n7 8 9\nv 0 1 0\nf 1 2 3\n")
    with pytest.raises(ff.FireflyError):
        ff.load_obj(str(bad))
    body = "v 0 0 0\nv 1 0 0\nv 0 1 0\nf 1 2 3\n"
    page = tmp_path / "page.obj"
    page.write_text("#" + "x" * (4096 - len(body) - 2) + "\n" + body)
    assert os.path.getsize(page) == 4096 and ff.load_obj(str(page)).shape == (1, 9)
    tail = tmp_path / "no_newline.obj"
    tail.write_text(body[:-1].replace("f 1 2 3", "f 1 2 -1"))
    assert ff.load_obj(str(tail)).shape == (1, 9)


def test_obj_reader_on_reference_assets(ff, meshes):
    for name in S.MESH_NAMES:
        got = ff.load_obj("/root/reference/sceneResources/%s.obj" % name)
        assert (bits(got) == bits(meshes[name])).all(), name


@pytest.fixture(scope="module")
def attrs():
    return S.load_mesh_attrs()


@pytest.mark.skipif(not os.path.isdir("/root/reference/sceneResources"), reason="reference tree not mounted")
def test_obj_attributes_on_reference_assets(ff, meshes, attrs):
    """Normals / uvs of the shipped meshes equal what the reference's own loader (tinyobjloader through LoadMesh's gather
    loop, oracle/ref_shim.cpp) produced for the committed fixture, bit for bit, for 1 and many parser threads."""
    for name in S.MESH_NAMES:
        for threads in (1, 0):
            m = ff.load_obj_ex("/root/reference/sceneResources/%s.obj" % name, threads=threads)
            assert (bits(m["xyz9"]) == bits(meshes[name])).all(), name
            assert (bits(m["normals9"]) == bits(attrs[name + "_n"])).all(), name
            assert (bits(m["uvs6"]) == bits(attrs[name + "_uv"])).all(), name
            assert m["has_normals"] == (name != "triangle") and m["has_uvs"] == (name != "triangle")   # sphere.obj: `vt` all zero


def _write_obj_full(path, xyz9, n9, uv6, face_mat, names, relative_every=0, mtl=None):
    """Indexed OBJ with v/vt/vn pools, `usemtl` runs, optional relative indices on every k-th face."""
    n = len(xyz9)
    with open(path, "w") as f:
        if mtl:
            f.write("mtllib %s\n" % mtl)
        nv = nt = nn = 0
        current = None
        for i in range(n):
            for k in range(3):
                f.write("v %.9g %.9g %.9g\n" % tuple(xyz9[i, 3 * k:3 * k + 3]))
                f.write("vn %.9g %.9g %.9g\n" % tuple(n9[i, 3 * k:3 * k + 3]))
                f.write("vt %.9g %.9g\n" % tuple(uv6[i, 2 * k:2 * k + 2]))
            nv += 3; nt += 3; nn += 3
            if face_mat[i] != current:
                current = face_mat[i]
                if current >= 0:
                    f.write("usemtl %s\n" % names[current])
            if relative_every and i % relative_every == 0:
                f.write("f -3/-3/-3 -2/-2/-2 -1/-1/-1\n")
            else:
                f.write("f %d/%d/%d %d/%d/%d %d/%d/%d\n" % (nv - 2, nt - 2, nn - 2, nv - 1, nt - 1, nn - 1, nv, nt, nn))


def test_obj_parallel_parse_and_materials(ff, meshes, attrs, tmp_path, monkeypatch):
    """The chunked parser gives the same arrays for any chunking (relative indices and `usemtl` runs cross chunk
    borders), and agrees with the reference's loader (tinyobjloader) on positions, normals, uvs and material ids."""
    name = "sphereBlender"
    xyz9, n9, uv6 = meshes[name], attrs[name + "_n"], attrs[name + "_uv"]
    n = len(xyz9)
    face_mat = ((np.arange(n) // 37) % 3).astype(np.int32)   # first use in order 0, 1, 2 = MTL order
    names = ["red", "mirror", "lamp"]
    (tmp_path / "lib.mtl").write_text(
        "newmtl red\nKd 0.9 0.1 0.1\nillum 2\n\nnewmtl mirror\nKd 0 0 0\nKs 0.8 0.8 0.9\nillum 3\n\n"
        "newmtl lamp\nKe 2 1.5 1\nKd 0 0 0\n")
    p = str(tmp_path / "full.obj")
    _write_obj_full(p, xyz9, n9, uv6, face_mat, names, relative_every=5, mtl="lib.mtl")
    base = ff.load_obj_ex(p, threads=1)
    assert (bits(base["xyz9"]) == bits(xyz9)).all() and (bits(base["normals9"]) == bits(n9)).all()
    assert (bits(base["uvs6"]) == bits(uv6)).all() and (base["face_material"] == face_mat).all()
    assert base["n_materials"] == 3 and base["has_normals"] and base["has_uvs"]
    monkeypatch.setenv("FF_OBJ_CHUNK_BYTES", "4096")   # ~100 chunks' worth of file, cut into up to `threads` chunks
    for threads in (2, 3, 7, 16):
        m = ff.load_obj_ex(p, threads=threads)
        for k in ("xyz9", "normals9", "uvs6", "face_material"):
            assert np.array_equal(m[k], base[k]), (threads, k)
    monkeypatch.delenv("FF_OBJ_CHUNK_BYTES")
    from oracle import oracle as O
    if O.Ref.available():
        import ctypes as C
        lib = C.CDLL(O.REF_SO)
        fn = lib.ref_load_obj_attrs
        fn.restype = C.c_long
        fn.argtypes = [C.c_char_p] + [C.c_void_p] * 4 + [C.c_long]
        r = [np.zeros((n, 9), np.float32), np.zeros((n, 9), np.float32), np.zeros((n, 6), np.float32), np.zeros(n, np.int32)]
        cwd = os.getcwd()
        os.chdir(tmp_path)   # tinyobjloader resolves mtllib against the working directory
        try:
            assert fn(b"full.obj", *[a.ctypes.data for a in r], n) == n
        finally:
            os.chdir(cwd)
        assert (bits(r[0]) == bits(base["xyz9"])).all() and (bits(r[1]) == bits(base["normals9"])).all()
        assert (bits(r[2]) == bits(base["uvs6"])).all() and (r[3] == base["face_material"]).all()
    # malformed input is reported with the file's line number, also from a later chunk
    bad = tmp_path / "bad2.obj"
    bad.write_text("v 0 0 0\n" * 2000 + "f 1 2 x\n")
    monkeypatch.setenv("FF_OBJ_CHUNK_BYTES", "1024")
    with pytest.raises(ff.FireflyError, match="line 2001"):
        ff.load_obj_ex(str(bad), threads=8)


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


def test_camera_controls_match_reference_bit_exact(ff, golden):
    """ff_camera_process_keyboard / _mouse against the reference's own Camera::ProcessKeyboard / ProcessMouseMovement /
    resetCamera (utilities.h:343-427), state and basis after every one of 6 x 64 operations (tools/make_golden_camera.py)."""
    g = golden("ref_camera_walk.npz")
    for k in range(6):
        ops, states, bases, up = g["ops%d" % k], g["states%d" % k], g["bases%d" % k], g["up%d" % k]
        cam = ff.Camera(states[0][:3], states[0][3], states[0][4], world_up=up)
        for i, op in enumerate(ops):
            if int(op[0]) == 0:
                cam.process_keyboard(int(op[1]))
            else:
                cam.process_mouse(op[1], op[2], True)
            assert (bits(cam.state()) == bits(states[i + 1])).all(), (k, i, op, cam.state(), states[i + 1])
            assert (bits(np.concatenate(cam.basis())) == bits(bases[i + 1])).all(), (k, i, op)
    cam = ff.Camera()
    with pytest.raises(ff.FireflyError):
        cam.process_keyboard(11)
    # unconstrained pitch is allowed to pass 89 degrees, constrained is clamped (utilities.h:393-399)
    assert ff.Camera(pitch=88.0).process_mouse(0, 100, False).state()[4] > 89.0
    assert ff.Camera(pitch=88.0).process_mouse(0, 100, True).state()[4] == np.float32(89.0)
    assert ff.Camera(pitch=-88.0).process_mouse(0, -100, True).state()[4] == np.float32(-89.0)


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


def test_bench_clock_sampler_keeps_only_the_timed_region(tmp_path):
    """bench.py starts nvidia-smi before the warm-up (its start-up used to perturb the first timed steps) and must then report
    only the rows sampled inside the timed region, with throttle reasons taken from those rows alone."""
    import datetime
    import importlib.util
    import time
    spec = importlib.util.spec_from_file_location("ff_bench", os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "bench.py"))
    bench = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(bench)

    class Done:
        def terminate(self): pass
        def wait(self, timeout=None): return 0
        def kill(self): pass

    now = time.time()

    def row(t, sm, power, thermal):
        ts = datetime.datetime.fromtimestamp(t).strftime("%Y/%m/%d %H:%M:%S.%f")[:-3]
        return "%s, %d, 1965, %.2f, 0x0000000000000000, Not Active, Not Active, %s, Not Active" % (ts, sm, power, thermal)

    path = tmp_path / "smi.csv"
    path.write_text("\n".join([row(now - 3.0, 300, 150.0, "Active"),      # warm-up: idle clocks and a throttle flag that must not count
                               row(now - 1.0, 1965, 480.0, "Not Active"), row(now - 0.5, 1950, 500.0, "Not Active"),
                               row(now + 2.0, 210, 140.0, "Not Active")]) + "\n")
    s = bench.ClockSampler(0)
    s.proc, s.path = Done(), str(path)
    s.t0, s.t1 = now - 1.5, now
    out = s.stop()
    assert out["samples"] == 2 and out["sm_mhz"] == 1957.5 and out["sm_max_mhz"] == 1965.0
    assert out["reasons"] == [] and out["power_w_max"] == 500.0
    # without marks (or unparsable timestamps) every row counts, as before
    path.write_text(row(now, 1965, 480.0, "Active") + "\n")
    s2 = bench.ClockSampler(0)
    s2.proc, s2.path = Done(), str(path)
    out2 = s2.stop()
    assert out2["samples"] == 1 and out2["reasons"] == ["sw_thermal_slowdown"]
