#!/usr/bin/env python3
"""Writes tests/golden/mesh_attrs.npz: the per-corner normals and uvs of the reference's six meshes, gathered by the
reference's own loader (tinyobjloader, vendored in /root/reference) through LoadMesh's gather loop re-hosted in
oracle/ref_shim.cpp (ref_load_obj_attrs; utilities.h:806-837), plus their per-face material ids.  Run in the build
container, where /root/reference is mounted; the fixture travels to the GPU box.

    python tools/make_golden_attrs.py
"""
import ctypes as C
import os
import sys

import numpy as np

ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
sys.path.insert(0, ROOT)
from gpupathtracer_b200 import scenes as S  # noqa: E402
from oracle import oracle as O  # noqa: E402


def main():
    O.build()
    lib = C.CDLL(O.REF_SO)
    fn = lib.ref_load_obj_attrs
    fn.restype = C.c_long
    fn.argtypes = [C.c_char_p] + [C.c_void_p] * 4 + [C.c_long]
    out = {}
    meshes = S.load_meshes()
    for k in S.MESH_NAMES:
        path = os.path.join(S.REFERENCE_OBJ_DIR, k + ".obj").encode()
        n = fn(path, None, None, None, None, 0)
        xyz9, n9, uv6, mid = np.zeros((n, 9), np.float32), np.zeros((n, 9), np.float32), np.zeros((n, 6), np.float32), np.zeros(n, np.int32)
        fn(path, xyz9.ctypes.data, n9.ctypes.data, uv6.ctypes.data, mid.ctypes.data, n)
        assert np.array_equal(xyz9, meshes[k]), k
        out[k + "_n"], out[k + "_uv"] = n9, uv6
        print(k, n, "normals" if n9.any() else "no normals", "uvs" if uv6.any() else "no uvs", "material ids", np.unique(mid))
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "mesh_attrs.npz"), **out)


if __name__ == "__main__":
    main()
