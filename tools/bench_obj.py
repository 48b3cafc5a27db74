#!/usr/bin/env python3
"""OBJ ingest throughput: writes a synthetic OBJ of `--tris` triangles (v / vt / vn / f lines, the
shape LoadMesh reads, utilities.h:781-840) and parses it with ff_load_obj_ex on 1 and on all host threads; when the
re-hosted reference is present (oracle/_ref, this container only) the same file also goes through the reference's own
loader (tinyobjloader + LoadMesh's gather loop, ref_load_obj) and the positions are compared bit for bit.  Host-only:
needs no GPU.  Prints one JSON line.

    python tools/bench_obj.py [--tris 2000000]
"""
import argparse
import ctypes as C
import json
import os
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
sys.path.insert(0, ROOT)
import gpupathtracer_b200 as ff  # noqa: E402


def write_obj(path, n, rng):
    """n triangles over a shared vertex grid (each vertex used by ~6 faces, like a real mesh), with vt and vn."""
    side = int(np.ceil(np.sqrt(n / 2))) + 1
    xs, ys = np.meshgrid(np.arange(side, dtype=np.float32), np.arange(side, dtype=np.float32))
    z = rng.uniform(-1, 1, xs.shape).astype(np.float32)
    with open(path, "w") as f:
        v = np.stack([xs.ravel(), ys.ravel(), z.ravel()], 1)
        f.write("".join("v %.6f %.6f %.6f\n" % tuple(r) for r in v))
        f.write("".join("vt %.6f %.6f\n" % (r[0] / side, r[1] / side) for r in v))
        f.write("".join("vn 0.000000 0.000000 1.000000\n" for _ in v))
        lines, k = [], 0
        for y in range(side - 1):
            for x in range(side - 1):
                a, b, c, d = y * side + x + 1, y * side + x + 2, (y + 1) * side + x + 1, (y + 1) * side + x + 2
                lines.append("f %d/%d/%d %d/%d/%d %d/%d/%d\n" % (a, a, a, b, b, b, c, c, c))
                lines.append("f %d/%d/%d %d/%d/%d %d/%d/%d\n" % (b, b, b, d, d, d, c, c, c))
                k += 2
                if k >= n:
                    break
            if k >= n:
                break
        f.write("".join(lines[:n]))
    return min(k, n)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--tris", type=int, default=2000000)
    a = ap.parse_args()
    rng = np.random.default_rng(3)
    with tempfile.TemporaryDirectory() as d:
        path = os.path.join(d, "big.obj")
        n = write_obj(path, a.tris, rng)
        size = os.path.getsize(path)
        out = dict(triangles=n, file_mb=size / 1e6, host_threads=os.cpu_count())
        one = ff.load_obj_ex(path, threads=1)
        many = ff.load_obj_ex(path, threads=0)
        assert len(one["xyz9"]) == n and (one["xyz9"].view(np.uint32) == many["xyz9"].view(np.uint32)).all()
        out.update(parse_s_1_thread=one["seconds"], parse_s_all_threads=many["seconds"],
                   mtris_s_1_thread=n / one["seconds"] / 1e6, mtris_s_all_threads=n / many["seconds"] / 1e6,
                   mb_s_all_threads=size / many["seconds"] / 1e6)
        ref_so = os.path.join(ROOT, "oracle", "_ref", "libfirefly_ref.so")
        if os.path.exists(ref_so):
            lib = C.CDLL(ref_so)
            fn = lib.ref_load_obj
            fn.restype = C.c_long
            fn.argtypes = [C.c_char_p, C.c_void_p, C.c_long]
            buf = np.zeros((n, 9), np.float32)
            t0 = time.perf_counter()
            got = fn(path.encode(), buf.ctypes.data, n)
            dt = time.perf_counter() - t0
            out.update(reference_loader_s=dt, reference_mtris_s=n / dt / 1e6, speedup_vs_reference=dt / many["seconds"],
                       positions_bit_identical_to_reference=bool(got == n and (buf.view(np.uint32) == many["xyz9"].view(np.uint32)).all()))
    print(json.dumps(out))


if __name__ == "__main__":
    main()
