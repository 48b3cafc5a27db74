"""Scene descriptions of the five BASELINE.json configs plus the reference's own main() scene.

A description is plain data -- a list of geometries (mesh name or triangle array, TRS, material, flip) and a camera
dict -- so the same description can be instantiated on the GPU library (`instantiate`) and, by the tests, on the CPU
oracle.  The reference hard-codes its one scene in main() (kernel.cu:227-266, 311-322); `reference_main` restates it.

Meshes are the reference's scene assets (sceneResources/*.obj).  /root/reference does not exist on the GPU box, so
`load_meshes` reads the triangle-position fixture made from them by tools/make_golden.py, falling back to the OBJ
files (through this library's own OBJ reader) where the reference tree is mounted.
"""
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
MESH_FIXTURE = os.path.join(_HERE, "..", "tests", "golden", "meshes.npz")
REFERENCE_OBJ_DIR = "/root/reference/sceneResources"
ATTR_FIXTURE = os.path.join(_HERE, "..", "tests", "golden", "mesh_attrs.npz")   # per-corner normals / uvs of the same meshes
MESH_NAMES = ["cube", "triangle", "rocketman", "sphere", "sphereBlender", "wahoo"]

DIFFUSE_GREY = dict(type=1, albedo=(0.8, 0.8, 0.8))
DIFFUSE_RED = dict(type=1, albedo=(1.0, 0.0, 0.0))        # kernel.cu:237-239
EMITTER_WHITE2 = dict(type=0, emissive=(1.0, 1.0, 1.0), intensity=2.0)   # kernel.cu:241-244


def load_meshes():
    if os.path.exists(MESH_FIXTURE):
        z = np.load(MESH_FIXTURE)
        return {k: z[k] for k in z.files}
    if os.path.isdir(REFERENCE_OBJ_DIR):
        import gpupathtracer_b200 as ff
        return {k: ff.load_obj(os.path.join(REFERENCE_OBJ_DIR, k + ".obj")) for k in MESH_NAMES}
    raise FileNotFoundError("no mesh fixture (%s) and no reference tree" % MESH_FIXTURE)


def load_mesh_attrs():
    """{name + "_n": [n,9] corner normals, name + "_uv": [n,6] corner uvs} as the reference's LoadMesh gathers them
    (utilities.h:822-824; zeros where the file has none).  Fixture made by tools/make_golden_attrs.py."""
    if os.path.exists(ATTR_FIXTURE):
        z = np.load(ATTR_FIXTURE)
        return {k: z[k] for k in z.files}
    if os.path.isdir(REFERENCE_OBJ_DIR):
        import gpupathtracer_b200 as ff
        out = {}
        for k in MESH_NAMES:
            m = ff.load_obj_ex(os.path.join(REFERENCE_OBJ_DIR, k + ".obj"))
            out[k + "_n"], out[k + "_uv"] = m["normals9"], m["uvs6"]
        return out
    raise FileNotFoundError("no attribute fixture (%s) and no reference tree" % ATTR_FIXTURE)


def camera(pos, yaw=-90.0, pitch=0.0, fov=70.0, width=256, height=256, near=0.1, far=1000.0):
    return dict(pos=tuple(pos), world_up=(0, 1, 0), yaw=yaw, pitch=pitch, fov=fov, width=width, height=height, near=near, far=far)


def geom(mesh, pos=(0, 0, 0), rot=(0, 0, 0), scale=(1, 1, 1), material=None, flip=False, plane=False, sphere=None,
         normals=None, uvs=None, face_materials=None):
    """mesh: name | [n,9] array | None; plane=True: GeometryType::PLANE; sphere=r: GeometryType::SPHERE of radius r.
    Shading attributes: normals / uvs = True (the named mesh's own, from the attribute fixture) or an
    [n,9] / [n,6] array; face_materials = (int[n] indices, [material dicts]), -1 = `material`."""
    return dict(mesh=mesh, pos=pos, rot=rot, scale=scale, material=material or DIFFUSE_GREY, flip=flip, plane=plane, sphere=sphere,
                normals=normals, uvs=uvs, face_materials=face_materials)


def flip_winding(xyz9):
    a = np.array(xyz9, np.float32).reshape(-1, 9).copy()
    a[:, 3:6], a[:, 6:9] = a[:, 6:9].copy(), a[:, 3:6].copy()
    return a


def config1(width=256, height=256):
    """cube.obj + triangle.obj, 256x256, 1 spp, closest hit."""
    return dict(name="C1", geoms=[geom("cube"), geom("triangle", pos=(-0.5, -0.5, 1.0))],
                camera=camera((0, 0, 4), fov=40.0, width=width, height=height), spp=1, max_bounces=1)


def _cube_faces(cube):
    """Splits the 12 cube triangles by the axis/sign of their face so the ceiling can be its own (emissive) geometry."""
    cube = np.asarray(cube, np.float32).reshape(-1, 9)
    top = np.array([np.allclose(t.reshape(3, 3)[:, 1], 0.5) for t in cube])
    return cube[top], cube[~top]


def config2(meshes, width=1024, height=1024, spp=64, max_bounces=4):
    """rocketman.obj inside cube.obj (the README box), ceiling emissive, 64 spp, 4 bounces diffuse."""
    top, rest = _cube_faces(meshes["cube"])
    return dict(name="C2",
                geoms=[geom(rest, scale=(5, 5, 5), flip=True),
                       geom(top, scale=(5, 5, 5), flip=True, material=EMITTER_WHITE2),
                       geom("rocketman", rot=(0, 90, 180), material=DIFFUSE_RED)],   # kernel.cu:229
                camera=camera((0, 0, 2.4), fov=70.0, width=width, height=height), spp=spp, max_bounces=max_bounces)


def config3(meshes, width=1920, height=1080, spp=256, max_bounces=8):
    """wahoo.obj + sphere.obj light + two-triangle floor, wavefront vs megakernel."""
    y = -0.07
    floor = np.array([[-60, y, 60, 60, y, 60, 60, y, -60], [-60, y, 60, 60, y, -60, -60, y, -60]], np.float32)
    return dict(name="C3",
                geoms=[geom("wahoo"), geom("sphere", pos=(0, 18, 0), material=dict(type=0, emissive=(1, 1, 1), intensity=5.0)),
                       geom(floor)],
                camera=camera((0, 8, 30), fov=40.0, width=width, height=height), spp=spp, max_bounces=max_bounces)


def reference_main(width=800, height=800):
    """The scene main() hard-codes (kernel.cu:227-266, 311-322): rocketman + four unit planes scaled x5."""
    return dict(name="main",
                geoms=[geom("rocketman", rot=(0, 90, 180), material=DIFFUSE_RED),
                       geom(None, pos=(0, 0, -2.5), scale=(5, 5, 5), plane=True, material=DIFFUSE_RED),
                       geom(None, pos=(0, 2.5, 0), rot=(90, 0, 0), scale=(5, 5, 5), plane=True, material=DIFFUSE_RED),
                       geom(None, pos=(0, -2.5, 0), rot=(90, 0, 0), scale=(5, 5, 5), plane=True, material=DIFFUSE_RED),
                       geom(None, pos=(0, 0, 2.5), scale=(5, 5, 5), plane=True, material=DIFFUSE_RED)],
                camera=camera((0, 0, 15), fov=70.0, width=width, height=height), spp=1, max_bounces=0)


def scaled_meshes(width=160, height=160):
    """Meshes under uniform and non-uniform scales below and above 1, plus a scaled plane.  The as-coded frame shows
    inverse(transpose(M)) * n WITHOUT re-normalising it (kernel.cu:117,183), so a scale changes the brightness and a scale
    below 1 pushes components past 255, where make_uchar3 wraps (kernel.cu:214): the scene that pins both."""
    return dict(name="scaled",
                geoms=[geom("rocketman", pos=(-1.6, 0, 0), rot=(0, 90, 180), scale=(0.5, 0.5, 0.5), material=DIFFUSE_RED),
                       geom("cube", pos=(1.6, 0.5, -1.0), rot=(20, 30, 10), scale=(2.0, 1.0, 3.0)),
                       geom("sphereBlender", pos=(0.2, -1.6, 1.0), rot=(0, 0, 35), scale=(0.7, 1.3, 0.7)),
                       geom("triangle", pos=(-0.5, 1.2, 2.0), scale=(1.5, 1.5, 1.5)),
                       geom(None, pos=(0, 0, -3.0), scale=(8, 8, 8), plane=True, material=DIFFUSE_RED)],
                camera=camera((0, 0, 8), fov=50.0, width=width, height=height), spp=1, max_bounces=0)


def showcase(meshes, width=512, height=512, spp=16, max_bounces=6):
    """Every declared BXDFTyp and GeometryType in one frame: the config-2 box with a glass
    rocketman, an analytic mirror sphere, an analytic emissive sphere and an analytic plane."""
    top, rest = _cube_faces(meshes["cube"])
    return dict(name="showcase",
                geoms=[geom(rest, scale=(5, 5, 5), flip=True),
                       geom(top, scale=(5, 5, 5), flip=True, material=EMITTER_WHITE2),
                       geom("rocketman", rot=(0, 90, 180), scale=(0.6, 0.6, 0.6), pos=(-0.9, 0, 0),
                            material=dict(type=3, ior=1.5, specular=(1, 1, 1), transmittance=(0.9, 1.0, 0.9))),
                       geom(None, sphere=0.8, pos=(1.1, -1.6, -0.6), scale=(1, 1.2, 1), rot=(0, 0, 20),
                            material=dict(type=2, specular=(0.9, 0.9, 0.95))),
                       geom(None, sphere=0.25, pos=(0.2, 1.2, 0.8), material=dict(type=0, emissive=(1.0, 0.6, 0.2), intensity=6.0)),
                       geom(None, plane=True, pos=(0, -2.2, 0), rot=(-90, 0, 0), scale=(3, 3, 3), material=dict(type=1, albedo=(0.2, 0.4, 0.8)))],
                camera=camera((0, 0, 2.4), fov=70.0, width=width, height=height), spp=spp, max_bounces=max_bounces)


def shaded(meshes, width=256, height=256, spp=8, max_bounces=5):
    """The shading inputs LoadMesh carries and the reference's tracer ignores, all in one frame:
    interpolated corner normals on diffuse, mirror and glass meshes (one under a non-uniform scale, so the normal matrix
    matters), uvs, and per-face materials on wahoo including emissive faces (they join the light list of NEE)."""
    top, rest = _cube_faces(meshes["cube"])
    nw = len(meshes["wahoo"])
    fm = ((np.arange(nw) // 97) % 4 - 1).astype(np.int32)
    face_mats = [dict(type=1, albedo=(0.2, 0.3, 0.9)), dict(type=2, specular=(0.9, 0.8, 0.6)),
                 dict(type=0, emissive=(1.0, 0.5, 0.1), intensity=3.0)]
    return dict(name="shaded",
                geoms=[geom(rest, scale=(5, 5, 5), flip=True),
                       geom(top, scale=(5, 5, 5), flip=True, material=EMITTER_WHITE2),
                       geom("wahoo", pos=(-0.8, -2.4, 0.0), scale=(0.2, 0.2, 0.2), normals=True, uvs=True,
                            material=dict(type=1, albedo=(0.8, 0.7, 0.6)), face_materials=(fm, face_mats)),
                       # sphere.obj carries per-corner (smooth) normals; sphereBlender.obj is faceted (one `vn` per face)
                       geom("sphere", pos=(1.2, -1.6, -0.3), scale=(0.35, 0.35, 0.35), normals=True,
                            material=dict(type=2, specular=(0.95, 0.95, 0.95))),
                       geom("sphere", pos=(1.0, 0.8, -1.0), scale=(0.24, 0.24, 0.24), normals=True,
                            material=dict(type=3, ior=1.45, transmittance=(1.0, 0.9, 0.9))),
                       geom("sphere", pos=(-1.3, 1.0, -1.0), rot=(20, 30, 0), scale=(0.24, 0.36, 0.2), normals=True,
                            material=dict(type=1, albedo=(0.2, 0.8, 0.3))),
                       geom("sphereBlender", pos=(0.0, 1.6, -1.5), scale=(0.4, 0.4, 0.4), normals=True, uvs=True,
                            material=dict(type=1, albedo=(0.7, 0.7, 0.2)))],
                camera=camera((0, 0, 2.4), fov=70.0, width=width, height=height), spp=spp, max_bounces=max_bounces)


def soup_scale(n):
    return float(np.float32(float(n) ** (-1.0 / 3.0)))


def config4_camera(width=3840, height=2160):
    return camera((0.5, 0.5, -1.5), yaw=90.0, fov=40.0, width=width, height=height)


def config5_light():
    """The one emitter of config 5: a two-triangle quad above the unit cube, facing down (front side towards the soup)."""
    y = 1.25
    quad = np.array([[-0.5, y, -0.5, 1.5, y, -0.5, 1.5, y, 1.5], [-0.5, y, -0.5, 1.5, y, 1.5, -0.5, y, 1.5]], np.float32)
    return geom(quad, material=dict(type=0, emissive=(1, 1, 1), intensity=4.0))


def config5(soup, width=7680, height=4320, spp=1024, max_bounces=4):
    """Synthetic N-triangle soup (20M in BASELINE.json) + one emissive quad, 8K, 1024 spp split across GPUs."""
    return dict(name="C5", geoms=[geom(soup), config5_light()], camera=config4_camera(width, height), spp=spp, max_bounces=max_bounces)


def resolve_mesh(g, meshes):
    m = g["mesh"]
    if isinstance(m, str):
        m = meshes[m]
    m = np.ascontiguousarray(m, np.float32).reshape(-1, 9)
    return flip_winding(m) if g["flip"] else m


def resolve_attrs(g, attrs=None):
    """(normals9 | None, uvs6 | None, (face_mat, materials) | None) of a mesh geometry, winding flip applied."""
    def pick(v, suffix, width):
        if v is None or v is False:
            return None
        if v is True:
            a = (attrs if attrs is not None else load_mesh_attrs())[g["mesh"] + suffix]
        else:
            a = v
        a = np.ascontiguousarray(a, np.float32).reshape(-1, 3 * width).copy()
        if g["flip"]:
            a[:, width:2 * width], a[:, 2 * width:] = a[:, 2 * width:].copy(), a[:, width:2 * width].copy()
        return a
    fm = g.get("face_materials")
    return pick(g.get("normals"), "_n", 3), pick(g.get("uvs"), "_uv", 2), (None if fm is None else (np.asarray(fm[0], np.int32), list(fm[1])))


def _apply_attrs(scene, gi, g, attrs):
    n9, uv6, fm = resolve_attrs(g, attrs)
    if n9 is not None:
        scene.set_normals(gi, n9)
    if uv6 is not None:
        scene.set_uvs(gi, uv6)
    if fm is not None:
        scene.set_face_materials(gi, fm[0], fm[1])


def instantiate(desc, ctx, meshes=None, builder=0, commit=True, attrs=None):
    """Builds the description on the GPU library; returns (Scene, Camera)."""
    import gpupathtracer_b200 as ff
    sc = ctx.scene()
    for g in desc["geoms"]:
        if g.get("sphere") is not None:
            sc.add_sphere(g["sphere"], g["pos"], g["rot"], g["scale"], g["material"])
        elif g["plane"]:
            sc.add_plane(g["pos"], g["rot"], g["scale"], g["material"])
        else:
            gi = sc.add_triangles(resolve_mesh(g, meshes), g["pos"], g["rot"], g["scale"], g["material"])
            _apply_attrs(sc, gi, g, attrs)
    if commit:
        sc.commit(builder)
    return sc, ff.Camera.from_dict(desc["camera"])


def instantiate_on(desc, scene, meshes=None, attrs=None):
    """Same on any object with add_mesh/add_plane (the test oracle's scene classes)."""
    for g in desc["geoms"]:
        if g.get("sphere") is not None:
            scene.add_sphere(g["sphere"], g["pos"], g["rot"], g["scale"], g["material"])
        elif g["plane"]:
            scene.add_plane(g["pos"], g["rot"], g["scale"], g["material"])
        else:
            gi = scene.add_mesh(resolve_mesh(g, meshes), g["pos"], g["rot"], g["scale"], g["material"])
            _apply_attrs(scene, gi, g, attrs)
    return scene


# --------------------------------------------------------------------------------------------- scene files
_TYPE_NAMES = {0: "emitter", 1: "diffuse", 2: "mirror", 3: "glass"}


def _f(v):
    return " ".join("%.9g" % float(np.float32(x)) for x in v)


def write_obj(path, xyz9):
    """Triangle positions -> a minimal OBJ (one 'v' per corner, one 'f' per triangle); float32 values round-trip exactly."""
    xyz9 = np.ascontiguousarray(xyz9, np.float32).reshape(-1, 9)
    with open(path, "w") as f:
        for t in xyz9:
            for k in range(3):
                f.write("v %s\n" % _f(t[3 * k:3 * k + 3]))
        for i in range(len(xyz9)):
            f.write("f %d %d %d\n" % (3 * i + 1, 3 * i + 2, 3 * i + 3))


def write_scene_file(desc, path, meshes=None, mesh_dir=None, render=None, inline_limit=64):
    """Writes a description as a scene file (format: csrc/ff_scenefile.cpp).  Named meshes and arrays of more than
    `inline_limit` triangles become OBJ files in `mesh_dir` (default: next to the scene file); small arrays become inline
    `triangles` blocks.  `render` overrides/extends the render line (dict: spp, bounces, seed, nee, aa, integrator, builder...)."""
    mesh_dir = mesh_dir or os.path.dirname(os.path.abspath(path))
    os.makedirs(mesh_dir, exist_ok=True)
    c = desc["camera"]
    lines = ["# %s -- written by gpupathtracer_b200.scenes.write_scene_file" % desc.get("name", "scene"),
             "camera pos %s up %s yaw %.9g pitch %.9g fov %.9g size %d %d near %.9g far %.9g" % (
                 _f(c["pos"]), _f(c.get("world_up", (0, 1, 0))), c["yaw"], c["pitch"], c["fov"], c["width"], c["height"], c["near"], c["far"])]
    mats = {}
    for gi, g in enumerate(desc["geoms"]):
        m = g["material"]
        key = repr(sorted((k, tuple(np.atleast_1d(v).tolist())) for k, v in m.items()))
        if key not in mats:
            mats[key] = "m%d" % len(mats)
            lines.append("material %s %s albedo %s emissive %s intensity %.9g specular %s ior %.9g transmittance %s" % (
                mats[key], _TYPE_NAMES[int(m.get("type", 1))], _f(m.get("albedo", (0.8, 0.8, 0.8))), _f(m.get("emissive", (0, 0, 0))),
                m.get("intensity", 0.0), _f(m.get("specular", (1, 1, 1))), m.get("ior", 1.5), _f(m.get("transmittance", (1, 1, 1)))))
        place = "pos %s rot %s scale %s material %s" % (_f(g["pos"]), _f(g["rot"]), _f(g["scale"]), mats[key])
        if g.get("sphere") is not None:
            lines.append("sphere %.9g %s" % (g["sphere"], place))
        elif g["plane"]:
            lines.append("plane %s" % place)
        else:
            tri = np.ascontiguousarray(meshes[g["mesh"]] if isinstance(g["mesh"], str) else g["mesh"], np.float32).reshape(-1, 9)
            flip = " flip" if g["flip"] else ""
            if isinstance(g["mesh"], str) or len(tri) > inline_limit:
                name = (g["mesh"] if isinstance(g["mesh"], str) else "geom%d" % gi) + ".obj"
                if not os.path.exists(os.path.join(mesh_dir, name)):
                    write_obj(os.path.join(mesh_dir, name), tri)
                lines.append("mesh %s %s%s" % (name, place, flip))
            else:
                lines.append("triangles %s%s" % (place, flip))
                lines += ["tri %s" % _f(t) for t in tri]
                lines.append("end")
    r = dict(spp=desc.get("spp", 1), bounces=desc.get("max_bounces", 4))
    r.update(render or {})
    lines.append("render " + " ".join("%s %s" % (k, _f(v) if isinstance(v, (tuple, list)) else v) for k, v in r.items()))
    with open(path, "w") as f:
        f.write("\n".join(lines) + "\n")
    return path
