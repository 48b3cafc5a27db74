"""Builds gpupathtracer_b200/libfirefly_b200.so in-tree with nvcc for sm_100a (cross-compiles without a GPU).

    python gpupathtracer_b200/build.py [--force]
"""
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OBJ = os.path.join(HERE, "build")
LIB = os.path.join(HERE, "libfirefly_b200.so")
SOURCES = ["ff_trace.cu", "ff_bvh.cu", "ff_sah.cu", "ff_render.cu", "ff_static.cu", "ff_api.cpp", "ff_group.cpp", "ff_io.cpp", "ff_obj.cpp", "ff_scenefile.cpp"]
CLI = os.path.join(HERE, "firefly_cli")   # headless front end, links against the C ABI only
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
FLAGS = ["-O3", "-std=c++17", "-lineinfo", "-Xcompiler", "-fPIC,-ffp-contract=off,-fno-fast-math,-Wall",
         "-Wno-deprecated-gpu-targets"] + os.environ.get("FF_EXTRA_NVCC_FLAGS", "").split()  # extra -D... for experiments


def _deps():
    return [os.path.join(CSRC, f) for f in os.listdir(CSRC)] + [os.path.join(HERE, "..", "include", "firefly.h"),
                                                                 os.path.abspath(__file__)]


def needs_build():
    if not os.path.exists(LIB) or not os.path.exists(CLI):
        return True
    t = os.path.getmtime(LIB)
    return any(os.path.getmtime(d) > t for d in _deps())


def _compile(src):
    out = os.path.join(OBJ, os.path.splitext(src)[0] + ".o")
    cmd = [NVCC] + ARCH + FLAGS + ["-c", os.path.join(CSRC, src), "-o", out]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("nvcc failed for %s:\n%s\n%s" % (src, r.stdout, r.stderr))
    return out


def build(force=False, verbose=False):
    if not force and not needs_build():
        return LIB
    os.makedirs(OBJ, exist_ok=True)
    with ThreadPoolExecutor(max_workers=len(SOURCES)) as ex:
        objs = list(ex.map(_compile, SOURCES))
    cmd = [NVCC] + ARCH + ["-shared", "-o", LIB] + objs + ["-Wno-deprecated-gpu-targets", "-ldl", "-lpthread"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("link failed:\n%s\n%s" % (r.stdout, r.stderr))
    # the CLI is plain C++ over include/firefly.h; rpath $ORIGIN so it finds the library next to it
    cmd = [os.environ.get("CXX", "g++"), "-O2", "-std=c++17", "-Wall", os.path.join(CSRC, "firefly_cli.cpp"), "-o", CLI,
           "-L" + HERE, "-lfirefly_b200", "-Wl,-rpath,$ORIGIN"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("CLI link failed:\n%s\n%s" % (r.stdout, r.stderr))
    if verbose:
        print("built", LIB, "and", CLI)
    return LIB


if __name__ == "__main__":
    build(force="--force" in sys.argv, verbose=True)
