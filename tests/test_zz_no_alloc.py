"""lmi_navigate_device_noalloc: the allocation-free navigation entry point (include/lmi_b200.h).

No counterpart in the reference.  The check is against the library's own
lmi_navigate -- predict_proba, model.py:214-229 -- on the same inputs: both run the same kernel, so the int64 bucket
lists must be identical and the probabilities bit-equal.  Collected last on purpose."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

MAX_ENTRIES = 1 << 20          # kNoAllocMaxEntries (csrc/lmi_api.cu)


def _model(rng, d_nav=96, hidden=(128, 64), n_categories=37):
    dims = [d_nav, *hidden, n_categories]
    w = [(rng.standard_normal((dims[i + 1], dims[i])) / np.sqrt(dims[i])).astype(np.float32) for i in range(len(dims) - 1)]
    b = [(0.1 * rng.standard_normal(dims[i + 1])).astype(np.float32) for i in range(len(dims) - 1)]
    return w, b


def _vp(t):
    return C.c_void_p(t.data_ptr())


def test_noalloc_navigation_matches_lmi_navigate_twice_in_a_row():
    import torch
    from lmi_b200.index import GpuLearnedIndex
    from lmi_b200._lib import check
    rng = np.random.default_rng(20231)
    n_cat = 37
    li = GpuLearnedIndex(n_cat)
    w, b = _model(rng, n_categories=n_cat)
    li.set_model(w, b)
    lib, h = li._lib, li._h
    dev = torch.device("cuda", li.device)
    stream = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)

    # three calls in a row on the one shared scratch: a large batch, then a smaller one with another n_top (the stale
    # tail of the first call stays in the array and must not show), then the first shape again
    shapes = [(3001, 7), (257, n_cat), (3001, 7)]
    xs = [rng.standard_normal((nq, 96)).astype(np.float32) for nq, _ in shapes]
    x_dev = [torch.from_numpy(x).to(dev) for x in xs]
    cls_dev = [torch.full((nq, nt), -7, dtype=torch.int64, device=dev) for nq, nt in shapes]
    prob_dev = [torch.full((nq, nt), -1.0, dtype=torch.float32, device=dev) for nq, nt in shapes]
    torch.cuda.synchronize(dev)

    free_before = None
    for i, (nq, nt) in enumerate(shapes):
        if i == 1:                                   # module and kernels are loaded by the first call
            free_before = torch.cuda.mem_get_info(dev)[0]
        check(lib.lmi_navigate_device_noalloc(h, _vp(x_dev[i]), nq, nt, _vp(cls_dev[i]), _vp(prob_dev[i]), stream))
    torch.cuda.synchronize(dev)
    assert torch.cuda.mem_get_info(dev)[0] == free_before, "device memory was allocated or freed by the no-alloc entry point"

    for i, (nq, nt) in enumerate(shapes):
        want_cls = np.empty((nq, nt), dtype=np.int64)
        want_prob = np.empty((nq, nt), dtype=np.float32)
        check(lib.lmi_navigate(h, C.c_void_p(xs[i].ctypes.data), nq, nt, C.c_void_p(want_cls.ctypes.data),
                               C.c_void_p(want_prob.ctypes.data)))
        got_cls, got_prob = cls_dev[i].cpu().numpy(), prob_dev[i].cpu().numpy()
        assert np.array_equal(got_cls, want_cls), f"call {i}: bucket lists differ from lmi_navigate"
        assert np.array_equal(got_prob, want_prob), f"call {i}: probabilities differ from lmi_navigate"
        assert got_cls.min() >= 0 and got_cls.max() < n_cat

    # probabilities are optional (NULL), like lmi_navigate
    only_cls = torch.empty((shapes[0][0], shapes[0][1]), dtype=torch.int64, device=dev)
    check(lib.lmi_navigate_device_noalloc(h, _vp(x_dev[0]), shapes[0][0], shapes[0][1], _vp(only_cls), C.c_void_p(0), stream))
    torch.cuda.synchronize(dev)
    assert torch.equal(only_cls, cls_dev[0])
    li.close()


def test_noalloc_navigation_rejects_sizes_beyond_the_static_scratch():
    import torch
    from lmi_b200.index import GpuLearnedIndex
    rng = np.random.default_rng(7)
    n_cat = 16
    li = GpuLearnedIndex(n_cat)
    li.set_model(*_model(rng, n_categories=n_cat))
    lib, h = li._lib, li._h
    dev = torch.device("cuda", li.device)
    x = torch.zeros((8, 96), dtype=torch.float32, device=dev)
    out = torch.zeros((8, n_cat), dtype=torch.int64, device=dev)
    nq_too_many = MAX_ENTRIES // n_cat + 1            # rejected before any pointer is touched
    rc = lib.lmi_navigate_device_noalloc(h, _vp(x), nq_too_many, n_cat, _vp(out), C.c_void_p(0), C.c_void_p(0))
    assert rc == -1, rc
    msg = lib.lmi_last_error().decode()
    assert "exceeds the static scratch" in msg and str(MAX_ENTRIES) in msg, msg
    # the largest accepted product is exactly the limit: the argument check lets it through (nq = 0 rows are a no-op)
    assert lib.lmi_navigate_device_noalloc(h, _vp(x), 0, n_cat, _vp(out), C.c_void_p(0), C.c_void_p(0)) == 0
    assert lib.lmi_navigate_device_noalloc(h, _vp(x), 8, n_cat + 1, _vp(out), C.c_void_p(0), C.c_void_p(0)) == -1
    li.close()
