#!/usr/bin/env python3
"""Writes tests/golden/ref_camera_walk.npz: sequences of camera controls (keyboard codes 0..10 and mouse deltas) applied to
the reference's own Camera struct (ProcessKeyboard / ProcessMouseMovement / resetCamera, utilities.h:343-427) through
oracle/ref_shim.cpp (ref_camera_walk), with the state {pos, yaw, pitch} and the basis after every operation.  Run in the
build container, where /root/reference is mounted; the fixture travels to the GPU box.

    python tools/make_golden_camera.py
"""
import ctypes as C
import os
import sys

import numpy as np

ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
sys.path.insert(0, ROOT)
from oracle import oracle as O  # noqa: E402


def walk(lib, world_up, state, ops):
    """Applies ops one at a time; returns (states [n+1,5], bases [n+1,9])."""
    fn = lib.ref_camera_walk
    fn.restype = None
    fn.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_long, C.c_void_p]
    up = np.asarray(world_up, np.float32)
    st = np.asarray(state, np.float32).copy()
    basis = np.zeros(9, np.float32)
    states, bases = [], []
    fn(up.ctypes.data, st.ctypes.data, None, 0, basis.ctypes.data)
    states.append(st.copy()); bases.append(basis.copy())
    for op in np.asarray(ops, np.float32).reshape(-1, 3):
        o = np.ascontiguousarray(op, np.float32)
        fn(up.ctypes.data, st.ctypes.data, o.ctypes.data, 1, basis.ctypes.data)
        states.append(st.copy()); bases.append(basis.copy())
    return np.stack(states), np.stack(bases)


def main():
    O.build()
    lib = C.CDLL(O.REF_SO)
    rng = np.random.default_rng(20261105)
    out = {}
    for k in range(6):
        n = 64
        ops = np.zeros((n, 3), np.float32)
        for i in range(n):
            if rng.random() < 0.7:
                ops[i] = (0, rng.integers(0, 11 if k == 5 else 10), 0)
            else:
                ops[i] = (1, np.float32(rng.uniform(-400, 400)), np.float32(rng.uniform(-400, 400)))
        state0 = np.array([*rng.uniform(-5, 5, 3), rng.uniform(-180, 180), rng.uniform(-80, 80)], np.float32)
        up = (0, 1, 0) if k != 4 else (0.1, 1.0, -0.2)
        states, bases = walk(lib, up, state0, ops)
        out["up%d" % k], out["ops%d" % k], out["states%d" % k], out["bases%d" % k] = np.asarray(up, np.float32), ops, states, bases
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "ref_camera_walk.npz"), **out)
    print("wrote ref_camera_walk.npz", {k: v.shape for k, v in out.items() if k.endswith("0")})


if __name__ == "__main__":
    main()
