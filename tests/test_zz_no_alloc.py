"""ff_hits_unpack_static (include/firefly.h): the allocation-free variant of ff_hits_unpack, whose device temporaries live in a
file-scope static __device__ array (csrc/ff_static.cu).  No counterpart in the reference.
Checked through the C ABI against the existing entry point, bit for bit (hit ids are integers, t is compared by
its bits), on two calls in a row that share the static array.  This file sorts last on purpose (pytest -x)."""
import ctypes as C

import numpy as np
import pytest

from conftest import bits

pytestmark = pytest.mark.gpu


def _same(a, b):
    return (a["geom"] == b["geom"]).all() and (a["tri"] == b["tri"]).all() and (bits(a["t"]) == bits(b["t"])).all()


def test_static_scratch_unpack_matches_the_allocating_entry_point(ff, ctx):
    sc = ctx.scene()
    sc.add_triangles(ctx.make_soup(30000))
    sc.add_triangles(ctx.make_soup(20000))        # a second geometry, so geometryIndex is not constant
    sc.commit()
    n1, n2 = ff.HITS_UNPACK_STATIC_MAX, 12345     # the maximum size first, then a short batch over the first one's leftovers
    d_rays, d_hits1, d_hits2 = ctx.alloc(n1 * 32), ctx.alloc(n1 * 16), ctx.alloc(n2 * 16)
    ctx.make_rays_uniform(0, n1, device_buffer=d_rays, host=False)
    sc.trace_closest_device(d_rays, n1, d_hits1, ff.TRACE_BVH8)
    ctx.make_rays_uniform(0, n2, salt=7, device_buffer=d_rays, host=False)
    sc.trace_closest_device(d_rays, n2, d_hits2, ff.TRACE_BVH8)
    want1, want2 = sc.hits_unpack(d_hits1, n1), sc.hits_unpack(d_hits2, n2)
    assert 0 < want1["hit"].sum() < n1 and len(np.unique(want1["geom"])) == 3  # hits on both geometries, and misses
    assert not _same({k: v[:n2] for k, v in want1.items()}, want2)              # the two batches really differ

    got1 = sc.hits_unpack_static(d_hits1, n1)
    got2 = sc.hits_unpack_static(d_hits2, n2)      # second call in a row: same array, previous contents still in it
    assert _same(got1, want1)
    assert _same(got2, want2)
    assert _same(sc.hits_unpack_static(d_hits1, n1), want1)  # and back again

    # outputs are optional, as in ff_hits_unpack
    t_only = np.empty(n2, np.float32)
    assert ff.lib.ff_hits_unpack_static(sc.h, d_hits2.ptr, n2, None, None, t_only.ctypes.data_as(C.c_void_p)) == 0
    assert (bits(t_only) == bits(want2["t"])).all()
    assert ff.lib.ff_hits_unpack_static(sc.h, d_hits2.ptr, 0, None, None, None) == 0

    # one record more than the static array holds: a clear error, nothing written
    with pytest.raises(ff.FireflyError) as e:
        sc.hits_unpack_static(d_hits1, ff.HITS_UNPACK_STATIC_MAX + 1)
    assert "FF_HITS_UNPACK_STATIC_MAX" in str(e.value) and str(ff.HITS_UNPACK_STATIC_MAX) in str(e.value)
