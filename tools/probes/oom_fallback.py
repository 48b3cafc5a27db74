import sys
sys.path.insert(0,'/root/repo')
import numpy as np, torch
import gpupathtracer_b200 as ff
from gpupathtracer_b200 import scenes as S
ctx=ff.Context(0); meshes=S.load_meshes()
desc=S.config2(meshes,1024,1024,64,4)
sc,cam=S.instantiate(desc,ctx,meshes,builder=1)
free,total=torch.cuda.mem_get_info()
hog=torch.empty(int(free-6e9),dtype=torch.uint8,device='cuda')   # leave ~6 GB: the 64M-path default (10 GB) cannot fit
img,st=sc.render(cam,64,4,want_stats=True,host=False)
print("rendered with", st["kernel_launches"], "launches, gpu_ms", round(st["gpu_ms"],1))
del hog
ref=sc.render(cam,64,4)
print("same image:", bool((np.asarray(sc.render(cam,64,4)).view(np.uint32)==np.asarray(ref).view(np.uint32)).all()))
