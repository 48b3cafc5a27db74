import os
import sys

import numpy as np
import pytest

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def _ensure_built():
    import importlib.util
    spec = importlib.util.spec_from_file_location("ff_build", os.path.join(ROOT, "gpupathtracer_b200", "build.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    mod.build()


@pytest.fixture(scope="session")
def ff():
    _ensure_built()
    import gpupathtracer_b200
    return gpupathtracer_b200


@pytest.fixture(scope="session")
def orc():
    """The CPU oracle (test infrastructure): oracle/libfirefly_oracle.so, compiled on demand with g++."""
    from oracle import oracle as O
    if not O.Oracle.available():
        O.build()
    return O.Oracle()


@pytest.fixture(scope="session")
def ref():
    """The re-hosted reference (oracle/_ref), present only where it was built from /root/reference."""
    from oracle import oracle as O
    if not O.Ref.available():
        if os.path.isdir("/root/reference/PathTracer"):
            O.build()
        else:
            pytest.skip("oracle/_ref not built and /root/reference not mounted")
    return O.Ref()


@pytest.fixture(scope="session")
def meshes():
    z = np.load(os.path.join(GOLDEN, "meshes.npz"))
    return {k: z[k] for k in z.files}


@pytest.fixture(scope="session")
def golden():
    def load(name):
        z = np.load(os.path.join(GOLDEN, name))
        return {k: z[k] for k in z.files}
    return load


@pytest.fixture(scope="session")
def ctx(ff):
    if ff.device_count() == 0:
        pytest.skip("no CUDA device")
    c = ff.Context(0)
    yield c
    c.close()


def bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


def assert_hits_equal(a, b, what=""):
    """Bit-exact closest-hit parity: geometryIndex, triangleIndex and the key t (inf on a miss)."""
    bad = np.nonzero((a["geom"] != b["geom"]) | (a["tri"] != b["tri"]))[0]
    assert bad.size == 0, "%s: %d hit-id mismatches, first at ray %d: (%d,%d) vs (%d,%d)" % (
        what, bad.size, bad[0], a["geom"][bad[0]], a["tri"][bad[0]], b["geom"][bad[0]], b["tri"][bad[0]])
    tb = np.nonzero(bits(a["t"]) != bits(b["t"]))[0]
    assert tb.size == 0, "%s: %d t mismatches, first at ray %d: %r vs %r" % (what, tb.size, tb[0], a["t"][tb[0]], b["t"][tb[0]])
