"""ulk_smooth_contours_static: the allocation-free variant of ulk_smooth_contours (scratch in a
file-scope device array, include/ultrack_b200.h).  It has no counterpart in the reference; the
checker is the existing entry point ulk_smooth_contours (itself pinned to scipy in
tests/test_kernel_logic_emu.py and tests/test_gpu_parity.py), and the bar is bit equality.

Collected last on purpose (file name).
"""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
_WARM = []


def _contours(shape, seed):
    rng = np.random.default_rng(seed)
    return ((rng.random(shape) > 0.8).astype(np.float32)
            * rng.choice([0.25, 0.5, 0.75, 1.0], size=shape).astype(np.float32))


def _both(lib, ct, sigma3, stream=None):
    """(existing entry point, static entry point) on copies of the same device volume."""
    import torch
    from ultrack_b200 import _lib
    shape3 = ct.shape if ct.ndim == 3 else (1,) + ct.shape
    shp, sg = (C.c_int64 * 3)(*shape3), (C.c_double * 3)(*sigma3)
    ref = torch.from_numpy(ct).cuda()
    got = torch.from_numpy(ct).cuda()
    torch.cuda.synchronize()
    sp = C.c_void_p(stream.cuda_stream) if stream is not None else None
    if not _WARM:
        # the first launch of a kernel loads its module (lazy loading) and the first call looks the
        # scratch address up: set-up, kept outside the "device memory unchanged" check below
        tiny = torch.zeros(64, device="cuda")
        _lib.check(lib, lib.ulk_smooth_contours_static(C.c_void_p(tiny.data_ptr()), (C.c_int64 * 3)(1, 8, 8),
                                                       (C.c_double * 3)(0.0, 1.0, 1.0), None))
        torch.cuda.synchronize()
        _WARM.append(1)
    _lib.check(lib, lib.ulk_smooth_contours(C.c_void_p(ref.data_ptr()), shp, sg, sp))
    ws0, free0 = _ws(lib), torch.cuda.mem_get_info()[0]
    _lib.check(lib, lib.ulk_smooth_contours_static(C.c_void_p(got.data_ptr()), shp, sg, sp))
    torch.cuda.synchronize()          # the static call only enqueues
    assert _ws(lib) == ws0, "the static entry point touched the library workspace"
    assert torch.cuda.mem_get_info()[0] >= free0, "device memory in use grew across the static call"
    return ref.cpu().numpy(), got.cpu().numpy()


def _ws(lib):
    out = (C.c_int64 * 3)()
    lib.ulk_workspace_stats(out)
    return tuple(out)


def test_static_smoothing_equals_allocating_entry_twice_in_a_row(cuda_lib):
    """Two calls back to back on the shared scratch (a large 3-D volume with three passes, then a
    smaller 2-D frame with two: the second call sees the first one's leftovers), each bit-equal
    to ulk_smooth_contours and within the scipy tolerance of the existing tests."""
    import scipy.ndimage as ndi
    for shape, sigma3, seed in (((24, 96, 80), (1.0, 2.0, 1.5), 0), ((200, 168), (0.0, 1.5, 1.5), 1),
                                ((24, 96, 80), (1.0, 2.0, 1.5), 2)):
        ct = _contours(shape, seed)
        ref, got = _both(cuda_lib, ct, sigma3)
        assert np.array_equal(ref.view(np.uint32), got.view(np.uint32))
        exp = ndi.gaussian_filter(ct, sigma3[3 - ct.ndim:])
        exp = exp / exp.max()
        assert np.abs(got.reshape(shape) - exp).max() <= 2e-7


def test_static_smoothing_on_a_side_stream_and_single_axis(cuda_lib):
    import torch
    st = torch.cuda.Stream()
    ct = _contours((16, 40, 48), 3)
    with torch.cuda.stream(st):
        ref, got = _both(cuda_lib, ct, (0.0, 0.0, 3.0), stream=st)     # one pass: result lands in the scratch, div copies back
    assert np.array_equal(ref.view(np.uint32), got.view(np.uint32))
    ref, got = _both(cuda_lib, np.zeros((8, 8), np.float32), (0.0, 1.0, 1.0))          # max 0: 0/0 like numpy
    assert np.isnan(ref).all() and np.isnan(got).all()


def test_static_smoothing_refuses_sizes_beyond_the_scratch(cuda_lib):
    import torch
    lim = cuda_lib.ulk_smooth_contours_static_max_voxels()
    assert lim == 4194304
    buf = torch.zeros(16, device="cuda")
    sg = (C.c_double * 3)(0.0, 1.0, 1.0)
    rc = cuda_lib.ulk_smooth_contours_static(C.c_void_p(buf.data_ptr()), (C.c_int64 * 3)(1, 2048, 2049), sg, None)
    assert rc == -12 and b"static scratch limit" in cuda_lib.ulk_last_error()
    rc = cuda_lib.ulk_smooth_contours_static(C.c_void_p(buf.data_ptr()), (C.c_int64 * 3)(1, 4, 4),
                                             (C.c_double * 3)(0.0, 32.0, 1.0), None)
    assert rc == -12 and b"window radius" in cuda_lib.ulk_last_error()
    # exactly at the limit is served
    ct = _contours((2048, 2048), 4)
    ref, got = _both(cuda_lib, ct, (0.0, 1.0, 1.0))
    assert np.array_equal(ref.view(np.uint32), got.view(np.uint32))


def test_static_smoothing_under_graph_capture(cuda_lib):
    """The point of the entry point: after one set-up call it can be captured into a CUDA graph
    (no allocation, no synchronisation, no host read-back) and replayed."""
    import torch
    from ultrack_b200 import _lib
    shape3 = (12, 64, 64)
    shp, sg = (C.c_int64 * 3)(*shape3), (C.c_double * 3)(1.0, 1.0, 2.0)
    a, b = _contours(shape3, 5), _contours(shape3, 6)
    ref_a, got_a = _both(cuda_lib, a, tuple(sg))             # set-up call (address lookup) outside the capture
    ref_b, _ = _both(cuda_lib, b, tuple(sg))
    buf = torch.from_numpy(a).cuda()
    work = torch.empty_like(buf)
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        work.copy_(buf)
        st = torch.cuda.current_stream()
        _lib.check(cuda_lib, cuda_lib.ulk_smooth_contours_static(C.c_void_p(work.data_ptr()), shp, sg, C.c_void_p(st.cuda_stream)))
    for src, ref in ((a, ref_a), (b, ref_b), (a, ref_a)):
        buf.copy_(torch.from_numpy(src))
        g.replay()
        torch.cuda.synchronize()
        assert np.array_equal(work.cpu().numpy().view(np.uint32), ref.view(np.uint32))
