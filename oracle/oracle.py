"""TEST INFRASTRUCTURE ONLY: ctypes loaders for the CPU oracle and the re-hosted reference.

Imported by tests/, tools/make_golden.py, __graft_entry__.smoke() and bench.py's cpu_baseline /
--impl reference legs only.  The product package (gpupathtracer_b200) never imports this module.

  * `Oracle`  -> oracle/libfirefly_oracle.so  (oracle/firefly_oracle.cpp, the restatement, "port")
  * `Ref`     -> oracle/_ref/libfirefly_ref.so (the reference's own text, compiled from
                 /root/reference by oracle/Makefile; present only where it was built, "reference")
Both expose the same scene/intersect/camera surface so tests can swap one for the other.
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ORACLE_SO = os.path.join(HERE, "libfirefly_oracle.so")
REF_SO = os.path.join(HERE, "_ref", "libfirefly_ref.so")

_f3 = C.c_float * 3


def build(force=False):
    """Compile the restatement (always) and the re-hosted reference (when /root/reference is mounted)."""
    args = ["make", "-C", HERE, "-s"] + (["-B"] if force else []) + ["all"]
    subprocess.run(args, check=True)


def _ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _vec3(v):
    return _f3(*[float(x) for x in v])


def mat8(material):
    """material: None | dict(type=0..3, albedo, emissive, intensity, specular, ior, transmittance) -> float32[15]
    (BXDF fields, utilities.h:77-88; type = BXDFTyp: 0 emitter, 1 diffuse, 2 mirror, 3 glass)."""
    if material is None:
        return None
    return np.array([material.get("type", 1), *material.get("albedo", (0.8, 0.8, 0.8)),
                     *material.get("emissive", (0, 0, 0)), material.get("intensity", 0.0),
                     *material.get("specular", (1, 1, 1)), material.get("ior", 1.5),
                     *material.get("transmittance", (1, 1, 1))], np.float32)


class _Base:
    prefix = ""
    so = ""

    def __init__(self):
        if not os.path.exists(self.so):
            raise FileNotFoundError(self.so)
        self.lib = C.CDLL(self.so)
        L, p = self.lib, self.prefix
        getattr(L, p + "scene_new").restype = C.c_void_p
        getattr(L, p + "scene_free").argtypes = [C.c_void_p]
        getattr(L, p + "time_intersect").restype = C.c_double
        self.threads = getattr(L, p + "hardware_threads")()

    def f(self, name):
        return getattr(self.lib, self.prefix + name)

    @classmethod
    def available(cls):
        return os.path.exists(cls.so)

    # -------- camera
    def camera_rays(self, cam, want_rays=True):
        """cam: dict(pos, world_up, yaw, pitch, fov, width, height, near, far) -> (rays[n,6], basis[3,3])."""
        w, h = int(cam["width"]), int(cam["height"])
        rays = np.empty((h * w, 6), np.float32) if want_rays else None
        basis = np.empty((3, 3), np.float32)
        self.f("camera_rays")(_vec3(cam["pos"]), _vec3(cam.get("world_up", (0, 1, 0))), C.c_float(cam["yaw"]),
                              C.c_float(cam["pitch"]), C.c_float(cam["fov"]), w, h, C.c_float(cam["near"]),
                              C.c_float(cam["far"]), _ptr(rays), _ptr(basis))
        return rays, basis

    def camera_inverse_matrices(self, cam):
        iv = np.empty(16, np.float32)
        ip = np.empty(16, np.float32)
        self.f("camera_inverse_matrices")(_vec3(cam["pos"]), _vec3(cam.get("world_up", (0, 1, 0))),
                                          C.c_float(cam["yaw"]), C.c_float(cam["pitch"]), C.c_float(cam["fov"]),
                                          int(cam["width"]), int(cam["height"]), C.c_float(cam["near"]),
                                          C.c_float(cam["far"]), _ptr(iv), _ptr(ip))
        return iv, ip

    def intersect_triangle(self, v9, o, d):
        v9 = np.ascontiguousarray(v9, np.float32).reshape(9)
        t = C.c_float(0)
        p = np.zeros(3, np.float32)
        n = np.zeros(3, np.float32)
        hit = self.f("intersect_triangle")(_ptr(v9), _vec3(o), _vec3(d), C.byref(t), _ptr(p), _ptr(n))
        return bool(hit), np.float32(t.value), p, n

    def cosine_sample_hemisphere(self, u1, u2):
        out = np.zeros(3, np.float32)
        self.f("cosine_sample_hemisphere")(C.c_float(u1), C.c_float(u2), _ptr(out))
        return out


class _SceneBase:
    def __init__(self, owner):
        self.o = owner
        self.h = C.c_void_p(owner.f("scene_new")())
        self.n_geoms = 0

    def close(self):
        if self.h:
            self.o.f("scene_free")(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def matrices(self, gi):
        m = np.empty(16, np.float32)
        inv = np.empty(16, np.float32)
        self.o.f("geometry_matrices")(self.h, int(gi), _ptr(m), _ptr(inv))
        return m, inv


class RefScene(_SceneBase):
    def add_mesh(self, xyz9, pos=(0, 0, 0), rot=(0, 0, 0), scale=(1, 1, 1), material=None):
        xyz9 = np.ascontiguousarray(xyz9, np.float32).reshape(-1, 9)
        self.n_geoms += 1
        return self.o.f("scene_add_mesh")(self.h, _ptr(xyz9), C.c_long(len(xyz9)), _vec3(pos), _vec3(rot), _vec3(scale))

    def add_plane(self, pos=(0, 0, 0), rot=(0, 0, 0), scale=(1, 1, 1), material=None):
        self.n_geoms += 1
        return self.o.f("scene_add_plane")(self.h, _vec3(pos), _vec3(rot), _vec3(scale))

    def intersect(self, rays, threads=None, mode=0):
        assert mode == 0, "the re-hosted reference only has its own object-space path"
        rays = np.ascontiguousarray(rays, np.float32).reshape(-1, 6)
        n = len(rays)
        out = dict(hit=np.empty(n, np.uint8), geom=np.empty(n, np.int32), tri=np.empty(n, np.int32),
                   t=np.empty(n, np.float32), point=np.empty((n, 3), np.float32), normal=np.empty((n, 3), np.float32))
        self.o.f("intersect")(self.h, _ptr(rays), C.c_long(n), _ptr(out["hit"]), _ptr(out["geom"]), _ptr(out["tri"]),
                              _ptr(out["t"]), _ptr(out["point"]), _ptr(out["normal"]), int(threads or self.o.threads))
        return out

    def time_intersect(self, rays, threads=None, reps=1, mode=0):
        rays = np.ascontiguousarray(rays, np.float32).reshape(-1, 6)
        return self.o.f("time_intersect")(self.h, _ptr(rays), C.c_long(len(rays)), int(threads or self.o.threads),
                                          int(reps), None, None, None)

    def render_as_coded(self, cam, threads=None, mode=0):
        w, h = int(cam["width"]), int(cam["height"])
        rgb = np.empty((h, w, 3), np.uint8)
        self.o.f("render_as_coded")(self.h, _vec3(cam["pos"]), _vec3(cam.get("world_up", (0, 1, 0))),
                                    C.c_float(cam["yaw"]), C.c_float(cam["pitch"]), C.c_float(cam["fov"]), w, h,
                                    C.c_float(cam["near"]), C.c_float(cam["far"]), _ptr(rgb),
                                    int(threads or self.o.threads))
        return rgb


class OracleScene(_SceneBase):
    def add_mesh(self, xyz9, pos=(0, 0, 0), rot=(0, 0, 0), scale=(1, 1, 1), material=None):
        xyz9 = np.ascontiguousarray(xyz9, np.float32).reshape(-1, 9)
        self.n_geoms += 1
        return self.o.f("scene_add_mesh")(self.h, _ptr(xyz9), C.c_long(len(xyz9)), _vec3(pos), _vec3(rot),
                                          _vec3(scale), _ptr(mat8(material)))

    def add_plane(self, pos=(0, 0, 0), rot=(0, 0, 0), scale=(1, 1, 1), material=None):
        self.n_geoms += 1
        return self.o.f("scene_add_plane")(self.h, _vec3(pos), _vec3(rot), _vec3(scale), _ptr(mat8(material)))

    def add_sphere(self, radius=1.0, pos=(0, 0, 0), rot=(0, 0, 0), scale=(1, 1, 1), material=None):
        self.n_geoms += 1
        return self.o.f("scene_add_sphere")(self.h, C.c_float(radius), _vec3(pos), _vec3(rot), _vec3(scale), _ptr(mat8(material)))

    # shading attributes; None clears
    def set_normals(self, gi, normals9):
        a = None if normals9 is None else np.ascontiguousarray(normals9, np.float32).reshape(-1, 9)
        self.o.f("scene_set_normals")(self.h, int(gi), _ptr(a))

    def set_uvs(self, gi, uvs6):
        a = None if uvs6 is None else np.ascontiguousarray(uvs6, np.float32).reshape(-1, 6)
        self.o.f("scene_set_uvs")(self.h, int(gi), _ptr(a))

    def set_face_materials(self, gi, face_material, materials):
        if face_material is None:
            self.o.f("scene_set_face_materials")(self.h, int(gi), None, None, 0)
            return
        fm = np.ascontiguousarray(face_material, np.int32)
        m = np.ascontiguousarray(np.stack([mat8(x) for x in materials]) if len(materials) else np.zeros((0, 15)), np.float32)
        self.o.f("scene_set_face_materials")(self.h, int(gi), _ptr(fm), _ptr(m), int(len(materials)))

    def intersect_uv(self, rays, threads=None):
        """Closest hit on the baked scene + the barycentrics of kernel.cu:80-81 (0 on misses / analytic primitives)."""
        rays = np.ascontiguousarray(rays, np.float32).reshape(-1, 6)
        n = len(rays)
        out = dict(geom=np.empty(n, np.int32), tri=np.empty(n, np.int32), u=np.empty(n, np.float32), v=np.empty(n, np.float32))
        self.o.f("intersect_uv")(self.h, _ptr(rays), C.c_long(n), _ptr(out["geom"]), _ptr(out["tri"]), _ptr(out["u"]),
                                 _ptr(out["v"]), int(threads or self.o.threads))
        return out

    def render_aov(self, cam, kind, threads=None):
        """Primary-ray feature buffer; kind 3 shading normal, 4 uv, 5 barycentrics, 6 albedo -> rgba[h,w,4]."""
        w, h = int(cam["width"]), int(cam["height"])
        out = np.empty((h, w, 4), np.float32)
        self.o.f("render_aov")(self.h, int(kind), _vec3(cam["pos"]), _vec3(cam.get("world_up", (0, 1, 0))),
                               C.c_float(cam["yaw"]), C.c_float(cam["pitch"]), C.c_float(cam["fov"]), w, h,
                               C.c_float(cam["near"]), C.c_float(cam["far"]), _ptr(out), int(threads or self.o.threads))
        return out

    def flat(self):
        self.o.f("flat_count").restype = C.c_long
        n = self.o.f("flat_count")(self.h)
        xyz9 = np.empty((n, 9), np.float32)
        geom = np.empty(n, np.int32)
        tri = np.empty(n, np.int32)
        self.o.f("flat_triangles")(self.h, _ptr(xyz9), _ptr(geom), _ptr(tri))
        return xyz9, geom, tri

    def intersect(self, rays, threads=None, mode=1):
        """mode 0: the reference's object-space path; mode 1: baked world-space triangles (the GPU contract)."""
        rays = np.ascontiguousarray(rays, np.float32).reshape(-1, 6)
        n = len(rays)
        out = dict(hit=np.empty(n, np.uint8), geom=np.empty(n, np.int32), tri=np.empty(n, np.int32),
                   t=np.empty(n, np.float32), point=np.empty((n, 3), np.float32), normal=np.empty((n, 3), np.float32))
        self.o.f("intersect")(self.h, int(mode), _ptr(rays), C.c_long(n), _ptr(out["hit"]), _ptr(out["geom"]),
                              _ptr(out["tri"]), _ptr(out["t"]), _ptr(out["point"]), _ptr(out["normal"]),
                              int(threads or self.o.threads))
        return out

    def time_intersect(self, rays, threads=None, reps=1, mode=1):
        rays = np.ascontiguousarray(rays, np.float32).reshape(-1, 6)
        return self.o.f("time_intersect")(self.h, int(mode), _ptr(rays), C.c_long(len(rays)),
                                          int(threads or self.o.threads), int(reps), None, None, None)

    def render_as_coded(self, cam, threads=None, mode=0):
        w, h = int(cam["width"]), int(cam["height"])
        rgb = np.empty((h, w, 3), np.uint8)
        self.o.f("render_as_coded")(self.h, int(mode), _vec3(cam["pos"]), _vec3(cam.get("world_up", (0, 1, 0))),
                                    C.c_float(cam["yaw"]), C.c_float(cam["pitch"]), C.c_float(cam["fov"]), w, h,
                                    C.c_float(cam["near"]), C.c_float(cam["far"]), _ptr(rgb),
                                    int(threads or self.o.threads))
        return rgb

    def render(self, cam, spp, max_bounces, seed=1234, first_sample=0, aa=0, background=(0, 0, 0), ray_eps=1e-4,
               threads=None, nee=0, want_shadow_rays=False):
        """As-intended path tracer. Returns (rgba[h,w,4] mean radiance, closest-hit rays traced[, shadow rays])."""
        w, h = int(cam["width"]), int(cam["height"])
        out = np.empty((h, w, 4), np.float32)
        fn = self.o.f("render")
        fn.restype = C.c_long
        shadow = C.c_long(0)
        n = fn(self.h, _vec3(cam["pos"]), _vec3(cam.get("world_up", (0, 1, 0))), C.c_float(cam["yaw"]),
               C.c_float(cam["pitch"]), C.c_float(cam["fov"]), w, h, C.c_float(cam["near"]), C.c_float(cam["far"]),
               int(spp), int(max_bounces), C.c_uint32(seed & 0xFFFFFFFF), C.c_uint32((seed >> 32) & 0xFFFFFFFF),
               int(first_sample), int(aa), _vec3(background), C.c_float(ray_eps), int(nee), _ptr(out), C.byref(shadow),
               int(threads or self.o.threads))
        if want_shadow_rays:
            return out, n, shadow.value
        return out, n


    def render_crop(self, cam, spp, max_bounces, x0, y0, cw, ch, seed=1234, first_sample=0, aa=0, background=(0, 0, 0),
                    ray_eps=1e-4, threads=None, nee=0):
        """Pixels [x0, x0+cw) x [y0, y0+ch) of render(cam, ...), bit for bit (paths are keyed by the full-frame pixel index)."""
        w, h = int(cam["width"]), int(cam["height"])
        assert 0 <= x0 and 0 <= y0 and x0 + cw <= w and y0 + ch <= h
        out = np.empty((ch, cw, 4), np.float32)
        fn = self.o.f("render_crop")
        fn.restype = C.c_long
        n = fn(self.h, _vec3(cam["pos"]), _vec3(cam.get("world_up", (0, 1, 0))), C.c_float(cam["yaw"]),
               C.c_float(cam["pitch"]), C.c_float(cam["fov"]), w, h, C.c_float(cam["near"]), C.c_float(cam["far"]),
               int(spp), int(max_bounces), C.c_uint32(seed & 0xFFFFFFFF), C.c_uint32((seed >> 32) & 0xFFFFFFFF),
               int(first_sample), int(aa), _vec3(background), C.c_float(ray_eps), int(nee), int(x0), int(y0), int(cw), int(ch),
               _ptr(out), int(threads or self.o.threads))
        return out, n


class Ref(_Base):
    prefix = "ref_"
    so = REF_SO

    def scene(self):
        return RefScene(self)

    def load_obj(self, path):
        fn = self.lib.ref_load_obj
        fn.restype = C.c_long
        fn.argtypes = [C.c_char_p, C.c_void_p, C.c_long]
        n = fn(path.encode(), None, 0)
        if n < 0:
            raise IOError(path)
        a = np.zeros((n, 9), np.float32)
        fn(path.encode(), a.ctypes.data, n)
        return a


class Oracle(_Base):
    prefix = "orc_"
    so = ORACLE_SO

    def scene(self):
        return OracleScene(self)

    def philox(self, ctr, key):
        c = np.array(ctr, np.uint32)
        k = np.array(key, np.uint32)
        out = np.zeros(4, np.uint32)
        self.lib.orc_philox4x32_10(_ptr(c), _ptr(k), _ptr(out))
        return out

    def sincos2pi(self, u):
        u = np.ascontiguousarray(u, np.float32)
        s = np.empty_like(u)
        c = np.empty_like(u)
        self.lib.orc_sincos2pi(_ptr(u), C.c_long(u.size), _ptr(s), _ptr(c))
        return s, c

    def onb(self, n):
        t = np.zeros(3, np.float32)
        b = np.zeros(3, np.float32)
        self.lib.orc_onb(_vec3(n), _ptr(t), _ptr(b))
        return t, b

    def make_soup(self, n, s=None):
        s = float(n) ** (-1.0 / 3.0) if s is None else s
        out = np.empty((n, 9), np.float32)
        self.lib.orc_make_soup(C.c_long(n), C.c_float(s), _ptr(out))
        return out

    def make_rays_uniform(self, first, n, salt=0):
        out = np.empty((n, 6), np.float32)
        self.lib.orc_make_rays_uniform(C.c_long(first), C.c_long(n), C.c_uint32(salt), _ptr(out))
        return out
