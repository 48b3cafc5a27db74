import sys
sys.path.insert(0, '/root/repo')
import gpupathtracer_b200 as ff
n = int(sys.argv[1]) if len(sys.argv) > 1 else 20000000
ctx = ff.Context(0)
soup = ctx.make_soup(n)
sc = ctx.scene(); sc.add_triangles(soup)
sc.commit(1); sc.commit(1)
print(n, sc.info()["build_ms"], sc.info()["n_nodes"], repr(sc.info()["sah_cost"]))
