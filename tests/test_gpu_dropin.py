"""The drop-in exercised with the reference's REAL argument types and call sequence:
1-based ``pd.DataFrame``s built exactly as search/search.py:71-72, 83-84 builds them, ``li.search(...,
use_threshold=True)`` as search.py:119-128 calls it, the single-bucket branch of search.py:129-140, the
``data_navigation['category']`` side effect (LearnedIndex.py:67), the 'category' column drop (:140-141), store reuse
without re-uploading, and the N2 pieces (feature-major upload, ragged out-of-order appends, on-disk cache)."""
import ctypes as C
import os

import numpy as np
import pandas as pd
import pytest

from conftest import load_golden, assert_ids_equal_up_to_ties, rows_with_small_buckets

pytestmark = pytest.mark.gpu


def _frames(g):
    nav = np.zeros((g["data_f16"].shape[0], g["queries_nav"].shape[1]), dtype=np.float32)
    data = pd.DataFrame(nav); data.index += 1              # search.py:71-72 (values unused at query time)
    data_search = pd.DataFrame(g["data_f16"]); data_search.index += 1      # search.py:83-84
    return data, data_search


def _li(g, **kw):
    from lmi_b200.index import GpuLearnedIndex
    return GpuLearnedIndex(g["n_categories"], d_nav=g["queries_nav"].shape[1], d=g["data_f16"].shape[1], **kw) \
        .set_model(g["weights"], g["biases"])


@pytest.mark.parametrize("tag", ["ref_mlp_n2400_c24", "ref_mlp5_n1600_c16"])
def test_search_with_reference_argument_types(tag):
    from oracle import lmi_oracle as O
    g = load_golden(tag)
    data, data_search = _frames(g)
    pc = g["pred_categories"]
    li = _li(g)
    nb, k = g["bucket_list"][0], g["k"]
    assert "category" not in data.columns
    d, a = li.search(data_navigation=data, queries_navigation=g["queries_nav"], data_search=data_search,
                     queries_search=g["queries_f16"], pred_categories=pc, n_buckets=nb, k=k, use_threshold=True)
    assert (data["category"].to_numpy() == pc).all()                       # LearnedIndex.py:67
    assert d.dtype == np.float64 and a.dtype == np.uint32 and d.shape == a.shape == (g["queries_f16"].shape[0], k)
    ok = ~rows_with_small_buckets(g, nb)
    assert_ids_equal_up_to_ties(a[ok], d[ok], g[f"clean_ids_nb{nb}"][ok], g[f"clean_dists_nb{nb}"][ok], what="frames vs reference")
    od, oa = O.search_intended(g["classes"], g["data_f16"], g["queries_f16"], pc, nb, k)
    assert_ids_equal_up_to_ties(a, d, oa, od, what="frames vs oracle")
    assert li.store_builds == 1
    # second call, same objects: nothing is rebuilt, identical answer; so for every stop-condition value (search.py:115)
    for nb2 in g["bucket_list"] + [nb]:
        d2, a2 = li.search(data, g["queries_nav"], data_search, g["queries_f16"], pc, n_buckets=nb2, k=k, use_threshold=True)
    assert li.store_builds == 1 and (a2 == a).all() and (d2 == d).all()
    # temporaries with the same content (df.values each call): content-compared, still no rebuild
    d3, a3 = li.search(data, g["queries_nav"], data_search, g["queries_f16"], pc.copy(), n_buckets=nb, k=k)
    assert li.store_builds == 1 and (a3 == a).all()
    # different pred_categories: rebuilt, and the answer follows the new buckets
    pc2 = (pc + 1) % g["n_categories"]
    d4, a4 = li.search(data, g["queries_nav"], data_search, g["queries_f16"], pc2, n_buckets=nb, k=k)
    assert li.store_builds == 2 and (data["category"].to_numpy() == pc2).all()
    od4, oa4 = O.search_intended(g["classes"], g["data_f16"], g["queries_f16"], pc2, nb, k)
    assert_ids_equal_up_to_ties(a4, d4, oa4, od4, what="after pred_categories changed")
    # a data_search frame that carries the 'category' column (LearnedIndex.py:140-141) gives the same answer
    ds2 = data_search.copy(); ds2["category"] = pc2
    d5, a5 = li.search(data, g["queries_nav"], ds2, g["queries_f16"], pc2, n_buckets=nb, k=k)
    assert (a5 == a4).all() and (d5 == d4).all()
    # in-place edit of the same objects is not seen until invalidate() (documented)
    pc2[:] = pc
    li.invalidate()
    d6, a6 = li.search(data, g["queries_nav"], data_search, g["queries_f16"], pc2, n_buckets=nb, k=k)
    assert (a6 == a).all() and (d6 == d).all()


def test_single_bucket_branch_of_run():
    """search.py:129-140: predict_proba, data['category'] = pred_categories, search_single(classes[:, 0])."""
    g = load_golden("ref_mlp_n2400_c24")
    data, data_search = _frames(g)
    li = _li(g)
    _, classes = li.model.predict_proba(g["queries_nav"])                 # search.py:130-132
    data["category"] = g["pred_categories"]                               # search.py:133
    d, a = li.search_single(data_navigation=data, data_search=data_search, queries_search=g["queries_f16"],
                            pred_categories=classes[:, 0], k=g["k"])
    ok = ~rows_with_small_buckets(g, 1) & (classes[:, 0] == g["classes"][:, 0])
    assert_ids_equal_up_to_ties(a[ok], d[ok], g["single_ids"][ok], g["single_dists"][ok], what="single-bucket branch")
    assert li.store_builds == 1
    d2, a2 = li.search_single(data, data_search, g["queries_f16"], classes[:, 0], k=g["k"])
    assert li.store_builds == 1 and (a2 == a).all()
    # the reference regroups by data_navigation['category'] on every call (LearnedIndex.py:143): a changed column is honoured
    from oracle import lmi_oracle as O
    pc2 = (g["pred_categories"] * 7 + 3) % g["n_categories"]
    data["category"] = pc2
    d3, a3 = li.search_single(data, data_search, g["queries_f16"], classes[:, 0], k=g["k"])
    assert li.store_builds == 2
    od, oa = O.search_intended(classes[:, :1], g["data_f16"], g["queries_f16"], pc2, 1, g["k"])
    assert_ids_equal_up_to_ties(a3, d3, oa, od, what="search_single after the category column changed")
    # reassigning an equal column (search.py:133 runs inside the loop): compared by content, no rebuild
    data["category"] = pc2.copy()
    li.search_single(data, data_search, g["queries_f16"], classes[:, 0], k=g["k"])
    assert li.store_builds == 2


def test_from_reference_shaped_object():
    import torch
    from lmi_b200.index import GpuLearnedIndex
    g = load_golden("ref_mlp5_n1600_c16")
    seq = torch.nn.Sequential(torch.nn.Linear(96, 256), torch.nn.ReLU(), torch.nn.Linear(256, 128), torch.nn.ReLU(),
                              torch.nn.Linear(128, g["n_categories"]))
    with torch.no_grad():
        for m, w, b in zip([x for x in seq if isinstance(x, torch.nn.Linear)], g["weights"], g["biases"]):
            m.weight.copy_(torch.from_numpy(w)); m.bias.copy_(torch.from_numpy(b))

    class Holder:                                   # li.LearnedIndex -> .model (NeuralNetwork) -> .model (Model) -> .layers
        pass
    model, nn_, li_ref = Holder(), Holder(), Holder()
    model.layers = seq; nn_.model = model; li_ref.model = nn_
    li = GpuLearnedIndex.from_reference(li_ref)
    assert li.n_categories == g["n_categories"] and li.d_nav == 96
    probs, classes = li.model.predict_proba(g["queries_nav"])
    np.testing.assert_allclose(probs, g["probs"], rtol=1e-4, atol=1e-7)


def test_feature_major_and_ragged_appends_build_the_same_store():
    """N2: the store does not depend on how the rows arrive -- one row-major block, the feature-major block of a
    DataFrame (uploaded as it lies), or ragged chunks appended out of order through the C-ABI."""
    from lmi_b200._lib import check
    g = load_golden("ref_mlp_n2400_c24")
    x, pc = np.ascontiguousarray(g["data_f16"]), g["pred_categories"].astype(np.int64)
    n, d = x.shape
    nb, k = 3, g["k"]

    li_a = _li(g); li_a.prepare(x, pc)
    ref = li_a._search_prepared(g["queries_nav"], g["queries_f16"], nb, k)
    bp = li_a.bucket_ptr()
    xf = np.asfortranarray(x)                                   # feature-major, like a DataFrame block
    li_b = _li(g); li_b.prepare(xf, pc)
    df = pd.DataFrame(x); df.index += 1
    li_c = _li(g); li_c.prepare(df, pc)
    li_d = _li(g)                                               # ragged, out of order, mixed layouts
    check(li_d._lib.lmi_store_begin(li_d._h, C.c_void_p(pc.ctypes.data), n, 1, 0))
    cuts = [0, 1, 64, 65, 700, 701, 1500, 2399, n]
    order = [5, 0, 7, 2, 4, 1, 6, 3]
    for j in order:
        lo, hi = cuts[j], cuts[j + 1]
        if j % 2 == 0:
            blk = np.ascontiguousarray(x[lo:hi])
            check(li_d._lib.lmi_store_append(li_d._h, C.c_void_p(blk.ctypes.data), lo, hi - lo, 0))
        else:                                                   # a column slice of the feature-major block: row_stride 1, col_stride n
            check(li_d._lib.lmi_store_append_strided(li_d._h, C.c_void_p(xf.ctypes.data + lo * 2), lo, hi - lo, 1, n, 0))
    check(li_d._lib.lmi_store_end(li_d._h))
    li_d._store_key = ("manual", n)
    for li in (li_b, li_c, li_d):
        assert (li.bucket_ptr() == bp).all()
        d_, a_ = li._search_prepared(g["queries_nav"], g["queries_f16"], nb, k)
        assert (a_ == ref[1]).all() and (d_ == ref[0]).all()
    # appending a row twice / missing rows is an error at lmi_store_end
    li_e = _li(g)
    check(li_e._lib.lmi_store_begin(li_e._h, C.c_void_p(pc.ctypes.data), n, 1, 0))
    check(li_e._lib.lmi_store_append(li_e._h, C.c_void_p(x.ctypes.data), 0, n - 5, 0))
    assert li_e._lib.lmi_store_end(li_e._h) != 0


def test_store_cache_file_roundtrip(tmp_path):
    g = load_golden("ref_mlp5_n1600_c16")
    x, pc = g["data_f16"], g["pred_categories"]
    nb, k = 3, g["k"]
    li = _li(g); li.prepare(x, pc, cache_dir=str(tmp_path))
    assert not li.store_from_cache
    files = [f for f in os.listdir(tmp_path) if f.startswith("lmi_store_")]
    assert len(files) == 1
    ref = li._search_prepared(g["queries_nav"], g["queries_f16"], nb, k)
    li2 = _li(g); li2.prepare(x, pc, cache_dir=str(tmp_path))
    assert li2.store_from_cache and (li2.bucket_ptr() == li.bucket_ptr()).all()
    got = li2._search_prepared(g["queries_nav"], g["queries_f16"], nb, k)
    assert (got[1] == ref[1]).all() and (got[0] == ref[0]).all()
    # other pred_categories -> another key -> another file, built not loaded
    pc2 = (pc + 1) % g["n_categories"]
    li3 = _li(g); li3.prepare(x, pc2, cache_dir=str(tmp_path))
    assert not li3.store_from_cache and len([f for f in os.listdir(tmp_path) if f.startswith("lmi_store_")]) == 2
    # a file whose key does not match is refused by the library
    path = os.path.join(tmp_path, files[0])
    assert li3._lib.lmi_store_load(li3._h, path.encode(), b"\x00" * 32) != 0
    # a sharded index does not accept an unsharded file
    from lmi_b200.index import GpuLearnedIndex
    li4 = GpuLearnedIndex(g["n_categories"], rank=1, world=2).set_model(g["weights"], g["biases"])
    key = li._store_key[0]
    assert li4._lib.lmi_store_load(li4._h, path.encode(), key) != 0


def test_pageable_and_pinned_queries_agree():
    import torch
    g = load_golden("ref_mlp_n2400_c24")
    li = _li(g); li.prepare(g["data_f16"], g["pred_categories"])
    nb, k = 2, g["k"]
    d1, a1 = li._search_prepared(g["queries_nav"], g["queries_f16"], nb, k)              # pageable numpy
    qn = torch.from_numpy(g["queries_nav"]).pin_memory(); qs = torch.from_numpy(g["queries_f16"]).pin_memory()
    d2, a2 = li._search_prepared(qn.numpy(), qs.numpy(), nb, k)                          # pinned
    assert (a1 == a2).all() and (d1 == d2).all()


def test_second_device_gets_its_own_kernel_attributes():
    """cudaFuncAttributeMaxDynamicSharedMemorySize is per device: an index on cuda:1 created after one on cuda:0
    (and a narrower MLP after a wider one) must launch (process-global attribute caches)."""
    import torch
    from oracle import lmi_oracle as O
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    g = load_golden("ref_mlp5_n1600_c16")
    outs = []
    for dev in (0, 1):
        li = _li(g, device=dev); li.prepare(g["data_f16"], g["pred_categories"])
        outs.append(li._search_prepared(g["queries_nav"], g["queries_f16"], 3, g["k"]))
    assert (outs[0][1] == outs[1][1]).all() and (outs[0][0] == outs[1][0]).all()


def test_narrow_mlp_after_wide_mlp_same_device():
    g5, g2 = load_golden("ref_mlp5_n1600_c16"), load_golden("ref_mlp_n2400_c24")
    li5 = _li(g5)
    p5, _ = li5.model.predict_proba(g5["queries_nav"])
    li2 = _li(g2)                                               # narrower stack: must not shrink li5's launch limit
    p2, _ = li2.model.predict_proba(g2["queries_nav"])
    p5b, _ = li5.model.predict_proba(g5["queries_nav"])
    assert (p5 == p5b).all()
    np.testing.assert_allclose(p2, g2["probs"], rtol=1e-4, atol=1e-7)
