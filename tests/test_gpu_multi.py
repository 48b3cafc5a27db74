"""Multi-GPU parity on real devices: one rank per GPU under torchrun, NCCL and peer-memory framebuffer
reduce against the single-GPU frame.  Skipped on a one-GPU box; the host logic is covered by tests/test_multigpu_cpu.py."""
import json
import os
import socket
import subprocess
import sys

import pytest

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
pytestmark = pytest.mark.gpu


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def test_sample_split_over_nccl_and_peer_memory(ff, ctx):
    n = ff.device_count()
    if n < 2:
        pytest.skip("needs >= 2 GPUs (have %d)" % n)
    world = 2 if n < 4 else 4
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(world), "--master-addr", "127.0.0.1",
           "--master-port", str(_free_port()), os.path.join(ROOT, "tools", "multigpu_check.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
    line = [l for l in r.stdout.splitlines() if l.startswith("{")][-1]
    out = json.loads(line)
    assert out["ok"] and out["peer_bit_identical_to_rank_order_sum"], out
    assert out["nccl_tiles_bit_identical"] and out["peer_tiles_bit_identical"], out


def test_group_of_one_is_the_single_gpu_render(ff, ctx, meshes):
    """The C-ABI group with world 1 (runs on a one-GPU box): ff_group_render == ff_render bit for bit, both splits, both reduce
    flags (nothing to reduce), host and device outputs; ff_group_sample_range == the Python statement."""
    import numpy as np
    from gpupathtracer_b200 import multigpu as M
    from gpupathtracer_b200 import scenes as S
    import ctypes as C
    desc = S.config2(meshes, 96, 80, 6, 3)
    sc, cam = S.instantiate(desc, ctx, meshes)
    full = sc.render(cam, 6, 3)
    g = M.Group.single(ctx)
    assert g.world == 1 and g.rank == 0
    for split in ("samples", "tiles"):
        for reduce in ("nccl", "peer"):
            img, st = g.render(sc, cam, 6, 3, split=split, reduce=reduce)
            assert (img.view(np.uint32) == full.view(np.uint32)).all(), (split, reduce)
            assert st["rays"] > 0 and st["reduce_ms"] >= 0
    d = ctx.alloc(96 * 80 * 16)
    img, _ = g.render(sc, cam, 6, 3, to_host=False, device_out=d)
    assert img is None and (d.download(np.float32, 96 * 80 * 4).view(np.uint32) == full.reshape(-1).view(np.uint32)).all()
    for spp in (0, 1, 7, 64):
        for world in (1, 2, 3, 8):
            for r in range(world):
                a, b = C.c_int(), C.c_int()
                assert ff.lib.ff_group_sample_range(r, world, spp, C.byref(a), C.byref(b)) == 0
                assert (a.value, b.value) == M.sample_range(r, world, spp)
    g.close()


def test_single_process_group_over_several_devices(ff, meshes):
    """ff_group_create: one process, one context + host thread per device, ncclCommInitAll inside; NCCL reduce and the fused
    peer reduce-scatter kernel against the one-GPU frame, and the tile split bit for bit."""
    import numpy as np
    from gpupathtracer_b200 import multigpu as M
    from gpupathtracer_b200 import scenes as S
    n = ff.device_count()
    if n < 2:
        pytest.skip("needs >= 2 GPUs (have %d)" % n)
    world = min(n, 4)
    g = M.Group.local(list(range(world)))
    assert g.world == world and len(g.contexts) == world
    desc = S.config2(meshes, 160, 128, 8 * world, 3)
    scenes = [S.instantiate(desc, c, meshes)[0] for c in g.contexts]
    cam = ff.Camera.from_dict(desc["camera"])
    full = scenes[0].render(cam, 8 * world, 3)
    partial = []
    for r, sc in enumerate(scenes):
        first, count = M.sample_range(r, world, 8 * world)
        buf = g.contexts[r].alloc(160 * 128 * 16)
        buf.zero()
        sc.render_accumulate(cam, buf, count, 3, first_sample=first)
        partial.append(buf.download(np.float32, 160 * 128 * 4).reshape(128, 160, 4))
    want = M.combine_sums(partial, 8 * world)
    for reduce in ("nccl", "peer"):
        img, st = g.render(scenes, cam, 8 * world, 3, split="samples", reduce=reduce)
        assert np.abs(img - full).max() <= 1e-5, reduce
        if reduce == "peer":
            assert (img.view(np.uint32) == want.view(np.uint32)).all()  # sums in rank order, bit for bit
        img, _ = g.render(scenes, cam, 8 * world, 3, split="tiles", reduce=reduce)
        assert (img[..., :3].view(np.uint32) == full[..., :3].view(np.uint32)).all(), reduce
    for sc in scenes:
        sc.close()
    g.close()


def test_headless_cli_on_several_gpus(ff, meshes, tmp_path):
    """firefly_cli --gpus N: plain C++ over the C ABI renders one frame on N devices (tile split: bit-identical to --gpus 1;
    sample split: within the fp32 re-association bound)."""
    import numpy as np
    from gpupathtracer_b200 import scenes as S
    n = ff.device_count()
    if n < 2:
        pytest.skip("needs >= 2 GPUs (have %d)" % n)
    desc = S.showcase(meshes, 96, 64, 8, 4)
    path = S.write_scene_file(desc, str(tmp_path / "show.ffscene"), meshes, render=dict(seed=5, nee=1))
    cli = os.path.join(os.path.dirname(ff.__file__), "firefly_cli")

    def run(*extra):
        out = str(tmp_path / ("f_%s.pfm" % "_".join(x.strip("-") for x in extra)))
        r = subprocess.run([cli, path, "--out", out] + list(extra), capture_output=True, text=True, timeout=600)
        assert r.returncode == 0, r.stderr
        st = json.loads(r.stdout.strip().splitlines()[-1])
        body = open(out, "rb").read().split(b"\n", 3)[3]
        return np.frombuffer(body, "<f4").reshape(64, 96, 3), st
    one, st1 = run("--gpus", "1")
    assert st1["gpus"] == 1
    tiles, st = run("--gpus", "2", "--split", "tiles", "--reduce", "nccl")
    assert st["gpus"] == 2 and (tiles.view(np.uint32) == one.view(np.uint32)).all()
    for reduce in ("nccl", "peer"):
        img, st = run("--gpus", "2", "--split", "samples", "--reduce", reduce)
        assert np.abs(img - one).max() <= 1e-5 * max(1.0, float(one.max())) and st["rays"] == st1["rays"]
