import os, sys, time
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))
import numpy as np, torch, torch.distributed as dist
rank, local = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
import gpupathtracer_b200 as ff
from gpupathtracer_b200 import scenes as S, multigpu as M
ctx = ff.Context(local); meshes = S.load_meshes()
sc, cam = S.instantiate(S.config2(meshes), ctx, meshes)
w, h = cam.width, cam.height
for it in range(4):
    torch.cuda.synchronize(); dist.barrier(); t0 = time.perf_counter()
    accum = torch.zeros((h, w, 4), dtype=torch.float32, device="cuda"); torch.cuda.synchronize(); t1 = time.perf_counter()
    st = sc.render_accumulate(cam, accum.data_ptr(), 64, 4, first_sample=rank * 64); t2 = time.perf_counter()
    dist.reduce(accum, dst=0, op=dist.ReduceOp.SUM); torch.cuda.synchronize(); t3 = time.perf_counter()
    if rank == 0:
        img = ctx.resolve(accum.data_ptr(), w, h, 128, host=True); t4 = time.perf_counter()
        print("iter %d zeros %.2f render %.2f reduce %.2f resolve->host %.2f ms" % (it, 1e3*(t1-t0), 1e3*(t2-t1), 1e3*(t3-t2), 1e3*(t4-t3)), file=sys.stderr)
dist.barrier(); dist.destroy_process_group()
