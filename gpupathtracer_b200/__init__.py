"""gpupathtracer_b200 -- Python host mirror of the C ABI in include/firefly.h (libfirefly_b200.so).

The product is the shared library; this module is a thin ctypes binding whose names follow the reference's host
surface (LoadMesh -> load_obj, Geometry ctor -> Scene.add_obj/add_triangles/add_plane, Camera, render).
There is no CPU fallback: importing works without a GPU (so the symbol table can be checked), but creating a
Context raises FireflyError when no CUDA device is present, and a missing library raises ImportError loudly.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("FF_LIB_OVERRIDE") or os.path.join(_HERE, "libfirefly_b200.so")  # override: kernel-variant experiments

if not os.path.exists(LIB_PATH):
    raise ImportError("%s is missing: run `python gpupathtracer_b200/build.py` (or __graft_entry__.build()); "
                      "there is no fallback implementation" % LIB_PATH)


class _LazyLibrary:
    """The shared library is mapped on FIRST USE, not on import: `import gpupathtracer_b200.scenes` (pure-Python scene
    descriptions, which bench.py's reference arm and the oracle tests share) must not pull the product library into a
    process that is only supposed to run the CPU oracle.  Every attribute access goes through the real CDLL."""
    _handle = None

    def _load(self):
        if _LazyLibrary._handle is None:
            h = C.CDLL(LIB_PATH)
            _bind(h)
            _LazyLibrary._handle = h
        return _LazyLibrary._handle

    def __getattr__(self, name):
        return getattr(self._load(), name)


lib = _LazyLibrary()

CAM_FORWARD, CAM_BACKWARD, CAM_LEFT, CAM_RIGHT, CAM_UP, CAM_DOWN, CAM_YAW_LEFT, CAM_YAW_RIGHT, CAM_PITCH_UP, CAM_PITCH_DOWN, CAM_RESET = range(11)
INTEGRATOR_WAVEFRONT, INTEGRATOR_MEGAKERNEL, INTEGRATOR_AS_CODED = 0, 1, 2
AOV_NORMAL, AOV_UV, AOV_BARY, AOV_ALBEDO = 3, 4, 5, 6  # primary-ray feature buffers through render(integrator=...)
OBJ_FLIP_WINDING, OBJ_NORMALS, OBJ_UVS, OBJ_MATERIALS = 1, 2, 4, 8
BUILDER_LBVH, BUILDER_SAH = 0, 1
TRACE_BVH8, TRACE_BRUTE, TRACE_BVH2 = 0, 1, 2
TRACE_COUNT = 0x100
TRACE_STATIC = 0x200  # one-thread-per-ray kernel instead of the persistent dynamic-fetch kernel
MAT_EMITTER, MAT_DIFFUSE, MAT_MIRROR, MAT_GLASS = 0, 1, 2, 3
SPLIT_SAMPLES, SPLIT_TILES = 0, 1
REDUCE_NCCL, REDUCE_PEER = 0, 1
GROUP_ID_BYTES = 128
HITS_UNPACK_STATIC_MAX = 1 << 20   # FF_HITS_UNPACK_STATIC_MAX


class FireflyError(RuntimeError):
    def __init__(self, code, message):
        super().__init__("firefly error %d: %s" % (code, message))
        self.code = code


class CCamera(C.Structure):
    _fields_ = [("pos", C.c_float * 3), ("world_up", C.c_float * 3), ("yaw_deg", C.c_float), ("pitch_deg", C.c_float),
                ("fov_deg", C.c_float), ("width", C.c_int), ("height", C.c_int), ("near_clip", C.c_float),
                ("far_clip", C.c_float)]


class CMaterial(C.Structure):
    _fields_ = [("type", C.c_int), ("albedo", C.c_float * 3), ("emissive", C.c_float * 3), ("intensity", C.c_float),
                ("specular", C.c_float * 3), ("ior", C.c_float), ("transmittance", C.c_float * 3)]


class CRenderParams(C.Structure):
    _fields_ = [("spp", C.c_int), ("max_bounces", C.c_int), ("seed", C.c_uint64), ("integrator", C.c_int), ("nee", C.c_int),
                ("first_sample", C.c_int), ("aa", C.c_int), ("background", C.c_float * 3), ("ray_eps", C.c_float),
                ("samples_per_pass", C.c_int), ("total_spp", C.c_int), ("tile_size", C.c_int), ("tile_rank", C.c_int),
                ("tile_world", C.c_int)]


class CStats(C.Structure):
    _fields_ = [("gpu_ms", C.c_double), ("trace_ms", C.c_double), ("h2d_ms", C.c_double), ("d2h_ms", C.c_double),
                ("rays", C.c_uint64), ("shadow_rays", C.c_uint64), ("paths", C.c_uint64), ("nodes_visited", C.c_uint64),
                ("tris_tested", C.c_uint64), ("kernel_launches", C.c_uint32), ("passes", C.c_uint32)]

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_}


class CSceneInfo(C.Structure):
    _fields_ = [("n_triangles", C.c_uint64), ("n_geometries", C.c_uint64), ("n_planes", C.c_uint64), ("n_nodes", C.c_uint64),
                ("node_bytes", C.c_uint64), ("triangle_bytes", C.c_uint64), ("depth", C.c_uint32), ("builder", C.c_uint32),
                ("bounds_lo", C.c_float * 3), ("bounds_hi", C.c_float * 3), ("build_ms", C.c_double), ("sah_cost", C.c_double),
                ("morton_bits", C.c_uint32), ("reserved", C.c_uint32)]

    def as_dict(self):
        d = {k: getattr(self, k) for k, _ in self._fields_}
        d["bounds_lo"] = list(self.bounds_lo)
        d["bounds_hi"] = list(self.bounds_hi)
        return d


class CSceneFile(C.Structure):
    _fields_ = [("camera", CCamera), ("params", CRenderParams), ("builder", C.c_int), ("n_geometries", C.c_int)]


def _bind(lib):
    """argtypes / restypes of every entry point of include/firefly.h (run once, when the library is first used)."""
    lib.ff_last_error.restype = C.c_char_p
    lib.ff_load_obj.restype = C.c_long
    lib.ff_load_obj.argtypes = [C.c_char_p, C.c_void_p, C.c_size_t]
    _vp, _sz = C.c_void_p, C.c_size_t
    lib.ff_context_create.argtypes = [C.POINTER(_vp), C.c_int]
    lib.ff_context_destroy.argtypes = [_vp]
    lib.ff_context_synchronize.argtypes = [_vp]
    lib.ff_timer_begin.argtypes = [_vp]
    lib.ff_scene_load_file.argtypes = [_vp, C.c_char_p, C.c_char_p, C.POINTER(_vp), C.POINTER(CSceneFile)]
    lib.ff_scene_file_validate.argtypes = [C.c_char_p, C.POINTER(CSceneFile)]
    lib.ff_context_release_scratch.argtypes = [_vp]
    lib.ff_timer_end.argtypes = [_vp, C.POINTER(C.c_double)]
    lib.ff_scene_create.argtypes = [_vp, C.POINTER(_vp)]
    lib.ff_scene_destroy.argtypes = [_vp]
    lib.ff_scene_add_obj.argtypes = [_vp, C.c_char_p, _vp, _vp, _vp, _vp, C.c_int, C.POINTER(C.c_int)]
    lib.ff_scene_add_triangles.argtypes = [_vp, _vp, _sz, _vp, _vp, _vp, _vp, C.POINTER(C.c_int)]
    lib.ff_scene_add_plane.argtypes = [_vp, _vp, _vp, _vp, _vp, C.POINTER(C.c_int)]
    lib.ff_scene_add_sphere.argtypes = [_vp, C.c_float, _vp, _vp, _vp, _vp, C.POINTER(C.c_int)]
    lib.ff_scene_set_normals.argtypes = [_vp, C.c_int, _vp]
    lib.ff_scene_set_uvs.argtypes = [_vp, C.c_int, _vp]
    lib.ff_scene_set_face_materials.argtypes = [_vp, C.c_int, _vp, _vp, C.c_int]
    lib.ff_scene_add_obj_ex.argtypes = [_vp, C.c_char_p, _vp, _vp, _vp, _vp, C.c_int, C.POINTER(C.c_int)]
    lib.ff_load_obj_ex.restype = C.c_long
    lib.ff_load_obj_ex.argtypes = [C.c_char_p, _vp, _vp, _vp, _vp, _sz, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int),
                                   C.POINTER(C.c_int), C.POINTER(C.c_double)]
    lib.ff_trace_closest_uv.argtypes = [_vp, _vp, _sz, C.c_int, _vp, _vp, _vp, _vp, _vp, C.POINTER(CStats)]
    lib.ff_scene_commit.argtypes = [_vp, C.c_int, C.POINTER(C.c_double)]
    lib.ff_scene_get_info.argtypes = [_vp, C.POINTER(CSceneInfo)]
    lib.ff_scene_get_flat.argtypes = [_vp, _vp, _vp, _vp, _sz]
    lib.ff_scene_get_matrices.argtypes = [_vp, C.c_int, _vp, _vp]
    lib.ff_trace_closest.argtypes = [_vp, _vp, _sz, C.c_int, _vp, _vp, _vp, C.POINTER(CStats)]
    lib.ff_trace_closest_device.argtypes = [_vp, _vp, _sz, C.c_int, _vp, C.POINTER(CStats)]
    lib.ff_trace_any.argtypes = [_vp, _vp, _vp, _sz, C.c_int, _vp, C.POINTER(CStats)]
    lib.ff_hits_unpack.argtypes = [_vp, _vp, _sz, _vp, _vp, _vp]
    lib.ff_hits_unpack_static.argtypes = [_vp, _vp, _sz, _vp, _vp, _vp]
    lib.ff_camera_rays.argtypes = [_vp, C.POINTER(CCamera), _vp, _vp]
    lib.ff_camera_basis.argtypes = [C.POINTER(CCamera), _vp, _vp, _vp]
    lib.ff_camera_inverse_matrices.argtypes = [C.POINTER(CCamera), _vp, _vp]
    lib.ff_render.argtypes = [_vp, C.POINTER(CCamera), C.POINTER(CRenderParams), _vp, _vp, C.POINTER(CStats)]
    lib.ff_render_u8.argtypes = [_vp, C.POINTER(CCamera), C.POINTER(CRenderParams), _vp, C.c_int, C.POINTER(CStats)]
    lib.ff_render_accumulate.argtypes = [_vp, C.POINTER(CCamera), C.POINTER(CRenderParams), _vp, C.POINTER(CStats)]
    lib.ff_resolve.argtypes = [_vp, _vp, C.c_int, C.c_int, C.c_int, _vp, _vp]
    lib.ff_reduce_resolve_peers.argtypes = [_vp, C.POINTER(_vp), C.c_int, C.c_int, C.c_int, C.c_int, _vp, _vp]
    lib.ff_make_soup.argtypes = [_vp, _sz, C.c_float, _vp]
    lib.ff_make_rays_uniform.argtypes = [_vp, C.c_uint64, _sz, C.c_uint32, _vp, _vp]
    lib.ff_make_rays_bounce.argtypes = [_vp, _vp, _vp, _sz, C.c_uint64, _vp, C.POINTER(_sz)]
    lib.ff_device_malloc.argtypes = [_vp, _sz, C.POINTER(_vp)]
    lib.ff_device_free.argtypes = [_vp, _vp]
    lib.ff_device_memset.argtypes = [_vp, _vp, C.c_int, _sz]
    lib.ff_memcpy_h2d.argtypes = [_vp, _vp, _vp, _sz]
    lib.ff_memcpy_d2h.argtypes = [_vp, _vp, _vp, _sz]
    lib.ff_host_alloc_pinned.argtypes = [_sz, C.POINTER(_vp)]
    lib.ff_host_free_pinned.argtypes = [_vp]
    lib.ff_ipc_get_handle.argtypes = [_vp, _vp, _vp]
    lib.ff_ipc_open_handle.argtypes = [_vp, _vp, C.POINTER(_vp)]
    lib.ff_ipc_close_handle.argtypes = [_vp, _vp]
    lib.ff_flush_l2.argtypes = [_vp]
    lib.ff_write_image.argtypes = [C.c_char_p, _vp, C.c_int, C.c_int]
    lib.ff_write_image_u8.argtypes = [C.c_char_p, _vp, C.c_int, C.c_int]
    lib.ff_philox4x32_10.argtypes = [_vp, _vp, _vp]
    lib.ff_tile_owned_pixels.restype = C.c_long
    lib.ff_tile_owned_pixels.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, _vp, _sz]
    lib.ff_camera_process_keyboard.argtypes = [C.POINTER(CCamera), C.c_int]
    lib.ff_camera_process_mouse.argtypes = [C.POINTER(CCamera), C.c_float, C.c_float, C.c_int]
    lib.ff_progressive_create.argtypes = [_vp, C.POINTER(_vp)]
    lib.ff_progressive_destroy.argtypes = [_vp]
    lib.ff_progressive_destroy.restype = None
    lib.ff_progressive_reset.argtypes = [_vp]
    lib.ff_progressive_frame.argtypes = [_vp, C.POINTER(CCamera), C.POINTER(CRenderParams), _vp, _vp, _vp, C.c_int, C.POINTER(C.c_int),
                                         C.POINTER(CStats)]
    lib.ff_debug_node_stats.argtypes = [_vp, _vp, _vp]
    lib.ff_philox4x32_10.restype = None
    lib.ff_context_destroy.restype = None
    lib.ff_scene_destroy.restype = None
    lib.ff_group_create.argtypes = [_vp, C.c_int, C.POINTER(_vp)]
    lib.ff_group_unique_id.argtypes = [_vp]
    lib.ff_group_create_rank.argtypes = [_vp, _vp, C.c_int, C.c_int, C.POINTER(_vp)]
    lib.ff_group_destroy.argtypes = [_vp]
    lib.ff_group_destroy.restype = None
    lib.ff_group_world.argtypes = [_vp]
    lib.ff_group_local_count.argtypes = [_vp]
    lib.ff_group_rank.argtypes = [_vp, C.c_int]
    lib.ff_group_context.argtypes = [_vp, C.c_int]
    lib.ff_group_context.restype = _vp
    lib.ff_group_sample_range.argtypes = [C.c_int, C.c_int, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int)]
    lib.ff_group_render.argtypes = [_vp, C.POINTER(_vp), C.POINTER(CCamera), C.POINTER(CRenderParams), C.c_int, C.c_int, _vp, _vp,
                                    C.POINTER(CStats), C.POINTER(C.c_double)]


def _check(rc):
    if rc != 0:
        raise FireflyError(rc, lib.ff_last_error().decode(errors="replace"))


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _f3(v):
    return None if v is None else np.ascontiguousarray(v, np.float32).reshape(3)


def device_count():
    return lib.ff_device_count()


def load_obj(path):
    """OBJ file -> float32 [n, 9] triangle positions (LoadMesh, utilities.h:781-840)."""
    n = lib.ff_load_obj(os.fsencode(path), None, 0)
    if n < 0:
        _check(int(n))
    out = np.zeros((n, 9), np.float32)
    if n:
        lib.ff_load_obj(os.fsencode(path), _p(out), n)
    return out


def load_obj_ex(path, threads=0):
    """OBJ file with every attribute LoadMesh keeps (utilities.h:806-837), parsed on `threads` host threads (0 = all):
    dict(xyz9[n,9], normals9[n,9], uvs6[n,6], face_material[n] (index in order of first `usemtl`, -1 none),
    has_normals, has_uvs, n_materials, seconds)."""
    hn, hu, nm, sec = C.c_int(0), C.c_int(0), C.c_int(0), C.c_double(0)
    n = lib.ff_load_obj_ex(os.fsencode(path), None, None, None, None, 0, int(threads), None, None, None, None)
    if n < 0:
        _check(int(n))
    xyz9, n9, uv6, fm = np.zeros((n, 9), np.float32), np.zeros((n, 9), np.float32), np.zeros((n, 6), np.float32), np.zeros(n, np.int32)
    n2 = lib.ff_load_obj_ex(os.fsencode(path), _p(xyz9), _p(n9), _p(uv6), _p(fm), n, int(threads), C.byref(hn), C.byref(hu),
                            C.byref(nm), C.byref(sec))
    if n2 < 0:
        _check(int(n2))
    return dict(xyz9=xyz9, normals9=n9, uvs6=uv6, face_material=fm, has_normals=bool(hn.value), has_uvs=bool(hu.value),
                n_materials=nm.value, seconds=sec.value)


def philox4x32_10(ctr, key):
    c = np.array(ctr, np.uint32)
    k = np.array(key, np.uint32)
    o = np.zeros(4, np.uint32)
    lib.ff_philox4x32_10(_p(c), _p(k), _p(o))
    return o


def material(type=MAT_DIFFUSE, albedo=(0.8, 0.8, 0.8), emissive=(0, 0, 0), intensity=0.0, specular=(1, 1, 1), ior=1.5,
             transmittance=(1, 1, 1)):
    """BXDF fields (utilities.h:77-88)."""
    m = CMaterial()
    m.type = int(type)
    m.albedo[:] = [float(x) for x in albedo]
    m.emissive[:] = [float(x) for x in emissive]
    m.intensity = float(intensity)
    m.specular[:] = [float(x) for x in specular]
    m.ior = float(ior)
    m.transmittance[:] = [float(x) for x in transmittance]
    return m


def _material_from(obj):
    if obj is None or isinstance(obj, CMaterial):
        return obj
    return material(obj.get("type", MAT_DIFFUSE), obj.get("albedo", (0.8, 0.8, 0.8)), obj.get("emissive", (0, 0, 0)),
                    obj.get("intensity", 0.0), obj.get("specular", (1, 1, 1)), obj.get("ior", 1.5), obj.get("transmittance", (1, 1, 1)))


class Camera:
    """Mirrors the Camera fields main() sets (kernel.cu:311-322)."""

    def __init__(self, pos=(0, 0, 15), yaw=-90.0, pitch=0.0, fov=70.0, width=800, height=800, near=0.1, far=1000.0,
                 world_up=(0, 1, 0)):
        self.c = CCamera()
        self.c.pos[:] = [float(x) for x in pos]
        self.c.world_up[:] = [float(x) for x in world_up]
        self.c.yaw_deg, self.c.pitch_deg, self.c.fov_deg = float(yaw), float(pitch), float(fov)
        self.c.width, self.c.height = int(width), int(height)
        self.c.near_clip, self.c.far_clip = float(near), float(far)

    @classmethod
    def from_dict(cls, d):
        return cls(d["pos"], d["yaw"], d["pitch"], d["fov"], d["width"], d["height"], d["near"], d["far"],
                   d.get("world_up", (0, 1, 0)))

    @property
    def width(self):
        return self.c.width

    @property
    def height(self):
        return self.c.height

    def basis(self):
        f, r, u = (np.zeros(3, np.float32) for _ in range(3))
        _check(lib.ff_camera_basis(C.byref(self.c), _p(f), _p(r), _p(u)))
        return f, r, u

    def inverse_matrices(self):
        iv, ip = np.zeros(16, np.float32), np.zeros(16, np.float32)
        _check(lib.ff_camera_inverse_matrices(C.byref(self.c), _p(iv), _p(ip)))
        return iv, ip

    # Camera::ProcessKeyboard / ProcessMouseMovement (utilities.h:343-404): direction = CAM_* (the reference's codes)
    def process_keyboard(self, direction):
        _check(lib.ff_camera_process_keyboard(C.byref(self.c), int(direction)))
        return self

    def process_mouse(self, x_delta, y_delta, constrain_pitch=True):
        _check(lib.ff_camera_process_mouse(C.byref(self.c), float(x_delta), float(y_delta), int(bool(constrain_pitch))))
        return self

    def state(self):
        """(pos xyz, yaw, pitch) as float32, the part of the camera the controls change."""
        return np.array([*self.c.pos, self.c.yaw_deg, self.c.pitch_deg], np.float32)


def render_params(spp=1, max_bounces=4, seed=1234, integrator=INTEGRATOR_WAVEFRONT, nee=0, first_sample=0, aa=0,
                  background=(0, 0, 0), ray_eps=1e-4, samples_per_pass=0, total_spp=0, tile_size=0, tile_rank=0, tile_world=0):
    p = CRenderParams()
    p.spp, p.max_bounces, p.seed, p.integrator, p.nee = int(spp), int(max_bounces), int(seed), int(integrator), int(nee)
    p.first_sample, p.aa = int(first_sample), int(aa)
    p.background[:] = [float(x) for x in background]
    p.ray_eps = float(ray_eps)
    p.samples_per_pass, p.total_spp = int(samples_per_pass), int(total_spp)
    p.tile_size, p.tile_rank, p.tile_world = int(tile_size), int(tile_rank), int(tile_world)
    return p


class DeviceBuffer:
    def __init__(self, ctx, nbytes):
        self.ctx, self.nbytes = ctx, int(nbytes)
        self.ptr = C.c_void_p()
        _check(lib.ff_device_malloc(ctx.h, self.nbytes, C.byref(self.ptr)))

    def upload(self, arr):
        arr = np.ascontiguousarray(arr)
        assert arr.nbytes <= self.nbytes
        _check(lib.ff_memcpy_h2d(self.ctx.h, self.ptr, _p(arr), arr.nbytes))

    def download(self, dtype, count):
        out = np.empty(count, dtype)
        assert out.nbytes <= self.nbytes
        _check(lib.ff_memcpy_d2h(self.ctx.h, _p(out), self.ptr, out.nbytes))
        return out

    def zero(self):
        _check(lib.ff_device_memset(self.ctx.h, self.ptr, 0, self.nbytes))

    def free(self):
        if self.ptr:
            lib.ff_device_free(self.ctx.h, self.ptr)
            self.ptr = C.c_void_p()

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


class Context:
    def __init__(self, device=0):
        self.h = C.c_void_p()
        _check(lib.ff_context_create(C.byref(self.h), int(device)))
        self.device = int(device)

    def close(self):
        if self.h:
            for _, p in self.__dict__.pop("_pinned", {}).values():
                lib.ff_host_free_pinned(p)
            if not getattr(self, "_borrowed", False):  # a context of multigpu.Group.local belongs to its ff_group
                lib.ff_context_destroy(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def scene(self):
        return Scene(self)

    def pinned_array(self, shape, dtype=np.float32, key=None):
        """A numpy array over page-locked host memory (cudaHostAlloc), cached per (shape, dtype, key) for the life of the
        context: device<->host copies to it are true DMA transfers.  Reused by later calls with the same shape and key."""
        key = (tuple(shape), np.dtype(dtype).str, key)
        cache = self.__dict__.setdefault("_pinned", {})
        if key not in cache:
            nbytes = int(np.prod(shape)) * np.dtype(dtype).itemsize
            p = C.c_void_p()
            _check(lib.ff_host_alloc_pinned(nbytes, C.byref(p)))
            buf = (C.c_char * nbytes).from_address(p.value)
            cache[key] = (np.frombuffer(buf, dtype=dtype).reshape(shape), p)
        return cache[key][0]

    def alloc(self, nbytes):
        return DeviceBuffer(self, nbytes)

    def synchronize(self):
        _check(lib.ff_context_synchronize(self.h))

    def flush_l2(self):
        _check(lib.ff_flush_l2(self.h))

    def release_scratch(self):
        _check(lib.ff_context_release_scratch(self.h))

    def timer_begin(self):
        """CUDA-event bracket on the context's stream (device clock); pair with timer_end() -> milliseconds."""
        _check(lib.ff_timer_begin(self.h))

    def timer_end(self):
        ms = C.c_double()
        _check(lib.ff_timer_end(self.h, C.byref(ms)))
        return ms.value

    def make_soup(self, n, s=None):
        s = float(n) ** (-1.0 / 3.0) if s is None else float(s)
        out = np.empty((n, 9), np.float32)
        _check(lib.ff_make_soup(self.h, n, C.c_float(s), _p(out)))
        return out

    def make_rays_uniform(self, first, n, salt=0, device_buffer=None, host=True):
        out = np.empty((n, 6), np.float32) if host else None
        _check(lib.ff_make_rays_uniform(self.h, int(first), int(n), int(salt), _p(out),
                                        device_buffer.ptr if device_buffer is not None else None))
        return out

    def camera_rays(self, cam, device_buffer=None, host=True):
        n = cam.width * cam.height
        out = np.empty((n, 6), np.float32) if host else None
        _check(lib.ff_camera_rays(self.h, C.byref(cam.c), _p(out), device_buffer.ptr if device_buffer is not None else None))
        return out

    def resolve(self, accum, width, height, total_spp, host=True, device_out=None, pinned=False):
        out = (self.pinned_array((height, width, 4)) if pinned else np.empty((height, width, 4), np.float32)) if host else None
        _check(lib.ff_resolve(self.h, accum.ptr if isinstance(accum, DeviceBuffer) else C.c_void_p(accum), width, height,
                              int(total_spp), C.c_void_p(device_out) if device_out is not None else None, _p(out)))
        return out

    def reduce_resolve_peers(self, ptrs, width, height, total_spp, host=True, device_out=None, pinned=False):
        arr = (C.c_void_p * len(ptrs))(*[C.c_void_p(int(p)) for p in ptrs])
        out = (self.pinned_array((height, width, 4)) if pinned else np.empty((height, width, 4), np.float32)) if host else None
        _check(lib.ff_reduce_resolve_peers(self.h, arr, len(ptrs), width, height, int(total_spp),
                                           C.c_void_p(device_out) if device_out is not None else None, _p(out)))
        return out


class Scene:
    def __init__(self, ctx):
        self.ctx = ctx
        self.h = C.c_void_p()
        _check(lib.ff_scene_create(ctx.h, C.byref(self.h)))

    def close(self):
        if self.h:
            lib.ff_scene_destroy(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- geometry (Geometry ctor, utilities.h:176-213)
    def add_obj(self, path, pos=(0, 0, 0), rot=(0, 0, 0), scale=(1, 1, 1), material=None, flip_winding=False):
        idx = C.c_int(-1)
        m = _material_from(material)
        _check(lib.ff_scene_add_obj(self.h, os.fsencode(path), _p(_f3(pos)), _p(_f3(rot)), _p(_f3(scale)),
                                    C.byref(m) if m is not None else None, int(flip_winding), C.byref(idx)))
        return idx.value

    def add_obj_ex(self, path, pos=(0, 0, 0), rot=(0, 0, 0), scale=(1, 1, 1), material=None, flags=0):
        """add_obj keeping the file's attributes: flags = OBJ_FLIP_WINDING | OBJ_NORMALS | OBJ_UVS | OBJ_MATERIALS."""
        idx = C.c_int(-1)
        m = _material_from(material)
        _check(lib.ff_scene_add_obj_ex(self.h, os.fsencode(path), _p(_f3(pos)), _p(_f3(rot)), _p(_f3(scale)),
                                       C.byref(m) if m is not None else None, int(flags), C.byref(idx)))
        return idx.value

    # ---- shading attributes (Triangle::m_n0..2 / m_uv0..2 utilities.h:163-170, material_ids utilities.h:837); None clears
    def set_normals(self, gi, normals9):
        a = None if normals9 is None else np.ascontiguousarray(normals9, np.float32).reshape(-1, 9)
        _check(lib.ff_scene_set_normals(self.h, int(gi), _p(a)))

    def set_uvs(self, gi, uvs6):
        a = None if uvs6 is None else np.ascontiguousarray(uvs6, np.float32).reshape(-1, 6)
        _check(lib.ff_scene_set_uvs(self.h, int(gi), _p(a)))

    def set_face_materials(self, gi, face_material, materials=()):
        if face_material is None:
            _check(lib.ff_scene_set_face_materials(self.h, int(gi), None, None, 0))
            return
        fm = np.ascontiguousarray(face_material, np.int32)
        arr = (CMaterial * max(1, len(materials)))(*[_material_from(m) for m in materials])
        _check(lib.ff_scene_set_face_materials(self.h, int(gi), _p(fm), C.cast(arr, C.c_void_p), len(materials)))

    def add_triangles(self, xyz9, pos=(0, 0, 0), rot=(0, 0, 0), scale=(1, 1, 1), material=None):
        xyz9 = np.ascontiguousarray(xyz9, np.float32).reshape(-1, 9)
        idx = C.c_int(-1)
        m = _material_from(material)
        _check(lib.ff_scene_add_triangles(self.h, _p(xyz9), len(xyz9), _p(_f3(pos)), _p(_f3(rot)), _p(_f3(scale)),
                                          C.byref(m) if m is not None else None, C.byref(idx)))
        return idx.value

    add_mesh = add_triangles

    def add_plane(self, pos=(0, 0, 0), rot=(0, 0, 0), scale=(1, 1, 1), material=None):
        idx = C.c_int(-1)
        m = _material_from(material)
        _check(lib.ff_scene_add_plane(self.h, _p(_f3(pos)), _p(_f3(rot)), _p(_f3(scale)),
                                      C.byref(m) if m is not None else None, C.byref(idx)))
        return idx.value

    def add_sphere(self, radius=1.0, pos=(0, 0, 0), rot=(0, 0, 0), scale=(1, 1, 1), material=None):
        """GeometryType::SPHERE (utilities.h:143,227): analytic sphere, front side only."""
        idx = C.c_int(-1)
        m = _material_from(material)
        _check(lib.ff_scene_add_sphere(self.h, C.c_float(radius), _p(_f3(pos)), _p(_f3(rot)), _p(_f3(scale)),
                                       C.byref(m) if m is not None else None, C.byref(idx)))
        return idx.value

    def commit(self, builder=BUILDER_LBVH):
        ms = C.c_double(0)
        _check(lib.ff_scene_commit(self.h, int(builder), C.byref(ms)))
        return ms.value

    def info(self):
        i = CSceneInfo()
        _check(lib.ff_scene_get_info(self.h, C.byref(i)))
        return i.as_dict()

    def flat(self):
        n = self.info()["n_triangles"]
        xyz9, geom, tri = np.empty((n, 9), np.float32), np.empty(n, np.int32), np.empty(n, np.int32)
        _check(lib.ff_scene_get_flat(self.h, _p(xyz9), _p(geom), _p(tri), n))
        return xyz9, geom, tri

    def matrices(self, gi):
        m, inv = np.empty(16, np.float32), np.empty(16, np.float32)
        _check(lib.ff_scene_get_matrices(self.h, int(gi), _p(m), _p(inv)))
        return m, inv

    # ---- closest hit (intersectRays, kernel.cu:127-176)
    def trace_closest(self, rays, mode=TRACE_BVH8, want_stats=False):
        rays = np.ascontiguousarray(rays, np.float32).reshape(-1, 6)
        n = len(rays)
        geom, tri, t = np.empty(n, np.int32), np.empty(n, np.int32), np.empty(n, np.float32)
        st = CStats()
        _check(lib.ff_trace_closest(self.h, _p(rays), n, int(mode), _p(geom), _p(tri), _p(t), C.byref(st)))
        out = dict(hit=(geom >= 0).astype(np.uint8), geom=geom, tri=tri, t=t)
        return (out, st.as_dict()) if want_stats else out

    def trace_closest_into(self, rays, geom, tri, t, mode=TRACE_BVH8):
        """ff_trace_closest with caller-owned host buffers (page-locked ones from Context.pinned_array make the copies DMA
        transfers that overlap the traversal); returns the stats of the call."""
        assert rays.dtype == np.float32 and rays.flags.c_contiguous and rays.ndim == 2 and rays.shape[1] == 6
        n = len(rays)
        assert geom.dtype == np.int32 and tri.dtype == np.int32 and t.dtype == np.float32 and len(geom) == len(tri) == len(t) == n
        st = CStats()
        _check(lib.ff_trace_closest(self.h, _p(rays), n, int(mode), _p(geom), _p(tri), _p(t), C.byref(st)))
        return st.as_dict()

    def trace_closest_uv(self, rays, mode=TRACE_BVH8):
        """trace_closest + the barycentrics of each hit (u, v weight v1, v2; kernel.cu:80-81)."""
        rays = np.ascontiguousarray(rays, np.float32).reshape(-1, 6)
        n = len(rays)
        geom, tri, t = np.empty(n, np.int32), np.empty(n, np.int32), np.empty(n, np.float32)
        u, v = np.empty(n, np.float32), np.empty(n, np.float32)
        _check(lib.ff_trace_closest_uv(self.h, _p(rays), n, int(mode), _p(geom), _p(tri), _p(t), _p(u), _p(v), None))
        return dict(hit=(geom >= 0).astype(np.uint8), geom=geom, tri=tri, t=t, u=u, v=v)

    def trace_closest_device(self, d_rays, n, d_hits, mode=TRACE_BVH8):
        st = CStats()
        _check(lib.ff_trace_closest_device(self.h, d_rays.ptr, int(n), int(mode), d_hits.ptr, C.byref(st)))
        return st.as_dict()

    def hits_unpack(self, d_hits, n):
        geom, tri, t = np.empty(n, np.int32), np.empty(n, np.int32), np.empty(n, np.float32)
        _check(lib.ff_hits_unpack(self.h, d_hits.ptr, int(n), _p(geom), _p(tri), _p(t)))
        return dict(hit=(geom >= 0).astype(np.uint8), geom=geom, tri=tri, t=t)

    def hits_unpack_static(self, d_hits, n):
        """hits_unpack through ff_hits_unpack_static: no device allocation, at most HITS_UNPACK_STATIC_MAX records."""
        geom, tri, t = np.empty(n, np.int32), np.empty(n, np.int32), np.empty(n, np.float32)
        _check(lib.ff_hits_unpack_static(self.h, d_hits.ptr, int(n), _p(geom), _p(tri), _p(t)))
        return dict(hit=(geom >= 0).astype(np.uint8), geom=geom, tri=tri, t=t)

    def trace_any(self, rays, tmax=None, mode=TRACE_BVH8):
        rays = np.ascontiguousarray(rays, np.float32).reshape(-1, 6)
        n = len(rays)
        tm = None if tmax is None else np.ascontiguousarray(tmax, np.float32)
        occ = np.empty(n, np.uint8)
        st = CStats()
        _check(lib.ff_trace_any(self.h, _p(rays), _p(tm), n, int(mode), _p(occ), C.byref(st)))
        return occ

    def make_rays_bounce(self, d_rays_in, d_hits_in, n, seed, d_rays_out):
        out_n = C.c_size_t(0)
        _check(lib.ff_make_rays_bounce(self.h, d_rays_in.ptr, d_hits_in.ptr, int(n), int(seed), d_rays_out.ptr, C.byref(out_n)))
        return out_n.value

    # ---- render (launchPathTrace, kernel.cu:218-221,342)
    def render(self, cam, spp=1, max_bounces=4, want_stats=False, host=True, device_out=None, pinned=False, **kw):
        """ff_render.  pinned=True returns the frame in the context's cached page-locked array (overwritten by the next
        such call) instead of a fresh pageable one."""
        p = render_params(spp, max_bounces, **kw)
        out = None
        if host:
            out = self.ctx.pinned_array((cam.height, cam.width, 4)) if pinned else np.empty((cam.height, cam.width, 4), np.float32)
        st = CStats()
        _check(lib.ff_render(self.h, C.byref(cam.c), C.byref(p), _p(out), device_out.ptr if device_out is not None else None,
                             C.byref(st)))
        return (out, st.as_dict()) if want_stats else out

    def render_u8(self, cam, spp=1, max_bounces=0, integrator=INTEGRATOR_AS_CODED, want_stats=False, **kw):
        p = render_params(spp, max_bounces, integrator=integrator, **kw)
        out = np.empty((cam.height, cam.width, 3), np.uint8)
        st = CStats()
        _check(lib.ff_render_u8(self.h, C.byref(cam.c), C.byref(p), _p(out), 0, C.byref(st)))
        return (out, st.as_dict()) if want_stats else out

    def render_accumulate(self, cam, d_accum, spp, max_bounces, **kw):
        p = render_params(spp, max_bounces, **kw)
        st = CStats()
        ptr = d_accum.ptr if isinstance(d_accum, DeviceBuffer) else C.c_void_p(int(d_accum))
        _check(lib.ff_render_accumulate(self.h, C.byref(cam.c), C.byref(p), ptr, C.byref(st)))
        return st.as_dict()


class Progressive:
    """ff_progressive: accumulation across frames that restarts when the camera or an image-changing parameter moves
    (the state the viewer loop, kernel.cu:331-359, carries once the tracer accumulates).  Close it before its scene."""

    def __init__(self, scene):
        self.scene = scene
        self.h = C.c_void_p()
        _check(lib.ff_progressive_create(scene.h, C.byref(self.h)))

    def close(self):
        if self.h:
            lib.ff_progressive_destroy(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def reset(self):
        _check(lib.ff_progressive_reset(self.h))

    def frame(self, cam, spp=1, max_bounces=4, want_u8=False, want_stats=False, **kw):
        """Adds `spp` samples; returns (float RGBA mean so far, total spp[, 8-bit RGB][, stats])."""
        p = render_params(spp, max_bounces, **kw)
        out = np.empty((cam.height, cam.width, 4), np.float32)
        u8 = np.empty((cam.height, cam.width, 3), np.uint8) if want_u8 else None
        total = C.c_int(0)
        st = CStats()
        _check(lib.ff_progressive_frame(self.h, C.byref(cam.c), C.byref(p), _p(out), None, _p(u8) if want_u8 else None, 0,
                                        C.byref(total), C.byref(st)))
        r = [out, total.value]
        if want_u8:
            r.append(u8)
        if want_stats:
            r.append(st.as_dict())
        return tuple(r)


def tile_owned_pixels(width, height, rank, world, tile_size=0):
    """Pixel ids rank `rank` of `world` renders under the interleaved-tile split (ff_tile_owned_pixels)."""
    n = lib.ff_tile_owned_pixels(int(width), int(height), int(tile_size), int(rank), int(world), None, 0)
    if n < 0:
        _check(int(n))
    out = np.empty(n, np.uint32)
    if n:
        lib.ff_tile_owned_pixels(int(width), int(height), int(tile_size), int(rank), int(world), _p(out), n)
    return out


def validate_scene_file(path):
    """Parses a scene description file without touching a device; returns the CSceneFile (camera, params, builder)."""
    d = CSceneFile()
    _check(lib.ff_scene_file_validate(os.fsencode(path), C.byref(d)))
    return d


def load_scene_file(ctx, path, asset_dir=None):
    """Scene description file -> (Scene (not committed), Camera, CSceneFile)."""
    d = CSceneFile()
    sc = Scene.__new__(Scene)
    sc.ctx, sc.h = ctx, C.c_void_p()
    _check(lib.ff_scene_load_file(ctx.h, os.fsencode(path), os.fsencode(asset_dir) if asset_dir else None, C.byref(sc.h), C.byref(d)))
    cam = Camera(pos=tuple(d.camera.pos), yaw=d.camera.yaw_deg, pitch=d.camera.pitch_deg, fov=d.camera.fov_deg, width=d.camera.width,
                 height=d.camera.height, near=d.camera.near_clip, far=d.camera.far_clip, world_up=tuple(d.camera.world_up))
    return sc, cam, d


def write_image(path, rgba):
    rgba = np.ascontiguousarray(rgba)
    if rgba.dtype == np.uint8:
        h, w = rgba.shape[:2]
        _check(lib.ff_write_image_u8(os.fsencode(path), _p(rgba), w, h))
    else:
        rgba = np.ascontiguousarray(rgba, np.float32)
        h, w = rgba.shape[:2]
        assert rgba.shape[2] == 4
        _check(lib.ff_write_image(os.fsencode(path), _p(rgba), w, h))
