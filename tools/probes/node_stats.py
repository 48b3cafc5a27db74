#!/usr/bin/env python3
"""Development probe: child-count histogram of the committed 8-wide hierarchy (how full the nodes are), for the 1M soup
and the C2 / C3 scenes, both builders.   python tools/probes/node_stats.py [--tris 1000000]"""
import argparse
import ctypes
import json
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))
import gpupathtracer_b200 as ff  # noqa: E402
from gpupathtracer_b200 import scenes as S  # noqa: E402


def stats(sc):
    ch = (ctypes.c_ulonglong * 9)()
    lt = (ctypes.c_ulonglong * 4)()
    ff.lib.ff_debug_node_stats.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p]
    rc = ff.lib.ff_debug_node_stats(sc.h, ch, lt)
    assert rc == 0, rc
    ch, lt = list(ch), list(lt)
    nodes = sum(ch)
    return dict(nodes=nodes, child_hist=ch, mean_children=sum(k * v for k, v in enumerate(ch)) / max(1, nodes), leaf_tri_hist=lt[1:],
                mean_leaf_tris=sum(k * v for k, v in enumerate(lt)) / max(1, sum(lt)))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--tris", type=int, default=1000000)
    a = ap.parse_args()
    ctx = ff.Context(0)
    soup = ctx.make_soup(a.tris)
    for b in (0, 1):
        sc = ctx.scene()
        sc.add_triangles(soup)
        sc.commit(b)
        print(json.dumps(dict(case="soup", builder=b, **stats(sc))), flush=True)
        sc.close()
    meshes = S.load_meshes()
    for name, desc in (("c2", S.config2(meshes)), ("c3", S.config3(meshes, spp=16))):
        for b in (0, 1):
            sc, cam = S.instantiate(desc, ctx, meshes, builder=b)
            print(json.dumps(dict(case=name, builder=b, **stats(sc))), flush=True)
            sc.close()


if __name__ == "__main__":
    main()
