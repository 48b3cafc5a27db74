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
