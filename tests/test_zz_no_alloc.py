"""l3d_collinearity_prealloc: the collinearity pass for embedders that may not allocate GPU
memory at run time (a service with a fixed memory budget, a latency-bound frame loop).
Its scratch space is a fixed array inside the library, so what has to be shown is that it
returns exactly what the allocating entry point returns -- also on a second call, when the
built-in array still holds the first call's leftovers -- and that it refuses inputs beyond
its limit instead of overrunning the array.  Device buffers are supplied by the test (torch),
as a real embedder would supply its own.

Collected last on purpose (file name), so a failure here cannot hide other tests under -x."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _call(lib, segs_dev, n, pairs_dev, capacity, count_dev):
    err = C.create_string_buffer(256)
    rc = lib.l3d_collinearity_prealloc(segs_dev.data_ptr() if segs_dev is not None else None, n, 2.0,
                                       pairs_dev.data_ptr(), capacity, count_dev.data_ptr(), None, err, 256)
    return rc, err.value.decode()


def test_prealloc_collinearity_matches_allocating_entry_point():
    import torch
    from line3d_b200 import capi, scenes
    lib = capi.lib()
    ctx = capi.Context(0)
    sc = scenes.wireframe_building(n_views=4, n_seg=900, seed=5)
    big, small = sc.views[0].segs, sc.views[1].segs[:300]          # second call is smaller: stale bits/counts would show
    capacity = 200000
    pairs_dev = torch.zeros(capacity * 3, dtype=torch.int32, device="cuda:0")
    count_dev = torch.zeros(1, dtype=torch.int32, device="cuda:0")
    for segs in (big, small, big):
        want = ctx.collinearity(segs, 2.0)
        segs_dev = torch.from_numpy(np.ascontiguousarray(segs, dtype=np.float32)).to("cuda:0")
        rc, msg = _call(lib, segs_dev, segs.shape[0], pairs_dev, capacity, count_dev)
        assert rc == 0, msg
        torch.cuda.synchronize()
        n = int(count_dev.item())
        assert n == want.shape[0] and n > 0
        got = np.frombuffer(pairs_dev[: n * 3].cpu().numpy().tobytes(), dtype=capi.COLLIN_DTYPE)
        assert np.array_equal(got["i"], want["i"]) and np.array_equal(got["j"], want["j"])
        assert np.array_equal(got["value"], want["value"])          # same kernels, same arithmetic: bit-exact
    # a result buffer that is too small is never overrun; the count still reports the true total
    guard = torch.full((64 * 3,), -7, dtype=torch.int32, device="cuda:0")
    segs_dev = torch.from_numpy(np.ascontiguousarray(big, dtype=np.float32)).to("cuda:0")
    rc, msg = _call(lib, segs_dev, big.shape[0], guard, 32, count_dev)
    assert rc == 0, msg
    torch.cuda.synchronize()
    assert int(count_dev.item()) == ctx.collinearity(big, 2.0).shape[0]
    assert bool((guard[32 * 3:] == -7).all())
    ctx.close()


def test_prealloc_collinearity_matches_the_oracle_directly():
    """Not only equal to its allocating sibling: equal to the CPU restatement of K_collinearity (cudawrapper.cu:476-535),
    pair for pair and bit for bit, on two calls in a row."""
    import torch
    from line3d_b200 import capi, scenes
    from oracle import oracle_py
    lib = capi.lib()
    sc = scenes.wireframe_building(n_views=3, n_seg=500, seed=9)
    pairs_dev = torch.zeros(100000 * 3, dtype=torch.int32, device="cuda:0")
    count_dev = torch.zeros(1, dtype=torch.int32, device="cuda:0")
    for segs in (sc.views[0].segs, sc.views[2].segs[:211]):
        rel = oracle_py.collinearity(segs, 2.0)
        iu, ju = np.nonzero(np.triu(rel, 1) > 0)                   # row-major = (i, j) order
        segs_dev = torch.from_numpy(np.ascontiguousarray(segs, dtype=np.float32)).to("cuda:0")
        rc, msg = _call(lib, segs_dev, segs.shape[0], pairs_dev, 100000, count_dev)
        assert rc == 0, msg
        torch.cuda.synchronize()
        n = int(count_dev.item())
        assert n == iu.shape[0] and n > 0
        got = np.frombuffer(pairs_dev[: n * 3].cpu().numpy().tobytes(), dtype=capi.COLLIN_DTYPE)
        assert np.array_equal(got["i"], iu) and np.array_equal(got["j"], ju)
        assert np.array_equal(got["value"], rel[iu, ju])


def test_prealloc_collinearity_refuses_misaligned_segments_and_foreign_streams():
    import torch
    from line3d_b200 import capi
    lib = capi.lib()
    buf = torch.zeros(4 * 64 + 4, dtype=torch.float32, device="cuda:0")
    pairs_dev = torch.zeros(30, dtype=torch.int32, device="cuda:0")
    count_dev = torch.zeros(1, dtype=torch.int32, device="cuda:0")
    err = C.create_string_buffer(256)
    rc = lib.l3d_collinearity_prealloc(buf.data_ptr() + 4, 64, 2.0, pairs_dev.data_ptr(), 10, count_dev.data_ptr(), None, err, 256)
    assert rc == 1 and b"aligned" in err.value
    if torch.cuda.device_count() > 1:                                 # a stream of another device than the current one
        with torch.cuda.device(1):
            other = torch.cuda.Stream()
        torch.cuda.set_device(0)
        rc = lib.l3d_collinearity_prealloc(buf.data_ptr(), 64, 2.0, pairs_dev.data_ptr(), 10, count_dev.data_ptr(), C.c_void_p(other.cuda_stream), err, 256)
        assert rc == 7 and b"device" in err.value
    s = torch.cuda.Stream()                                           # a stream of the current device is fine
    rc = lib.l3d_collinearity_prealloc(buf.data_ptr(), 64, 2.0, pairs_dev.data_ptr(), 10, count_dev.data_ptr(), C.c_void_p(s.cuda_stream), err, 256)
    assert rc == 0, err.value
    torch.cuda.synchronize()


def test_prealloc_collinearity_rejects_oversized_input():
    import torch
    from line3d_b200 import capi
    lib = capi.lib()
    segs_dev = torch.zeros(4097 * 4, dtype=torch.float32, device="cuda:0")
    pairs_dev = torch.zeros(30, dtype=torch.int32, device="cuda:0")
    count_dev = torch.zeros(1, dtype=torch.int32, device="cuda:0")
    rc, msg = _call(lib, segs_dev, 4097, pairs_dev, 10, count_dev)
    assert rc == 2 and "4096" in msg
    rc, msg = _call(lib, segs_dev, 4096, pairs_dev, 10, count_dev)    # exactly at the limit is accepted
    assert rc == 0, msg
    torch.cuda.synchronize()
    assert int(count_dev.item()) == 0                                  # all-zero segments: nothing collinear
