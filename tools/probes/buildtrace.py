import sys
sys.path.insert(0,'/root/repo')
import gpupathtracer_b200 as ff
ctx=ff.Context(0)
for n in (1000000, 20000000):
    soup=ctx.make_soup(n)
    for b in (0,1):
        sc=ctx.scene(); sc.add_triangles(soup)
        sc.commit(b); print("---- n",n,"builder",b, file=sys.stderr, flush=True); sc.commit(b)
        print(n,b,sc.info()["build_ms"], file=sys.stderr, flush=True)
        sc.close()
