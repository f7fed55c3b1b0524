"""Live check of the oracle against the UNMODIFIED reference on inputs no fixture froze.

Runs wherever ``/root/reference`` exists (the build container); skipped elsewhere (the GPU box: the reference does
not travel, ``tests/golden`` does).  The reference hot path is imported in place through ``oracle/ref_shims.py``
(stub faiss / h5py, the numpy-2 ``np.ogrid`` shim; no reference file is edited) and called exactly as
``search/search.py:71-72, 83-84, 115-140`` calls it: 1-based DataFrames, ``search(..., use_threshold=...)``,
``predict_proba`` + ``search_single`` for one bucket.  Fresh seeds per parametrisation; weights are random (parity
needs the same weights on both sides, not trained ones).
"""
import numpy as np
import pytest

from oracle import lmi_oracle as O
from oracle import ref_shims
from conftest import assert_ids_equal_up_to_ties

pytestmark = pytest.mark.skipif(not ref_shims.reference_available(), reason="/root/reference not present")

K = 10


def _case(seed, n, nq, C, model_type):
    import pandas as pd
    rng = np.random.default_rng(seed)
    dn, d = 96, 768
    centres = rng.standard_normal((C * 3, d)).astype(np.float32)
    data = centres[rng.integers(0, C * 3, n)] + 0.8 * rng.standard_normal((n, d)).astype(np.float32)
    qs = centres[rng.integers(0, C * 3, nq)] + 0.8 * rng.standard_normal((nq, d)).astype(np.float32)
    data16, q16 = data.astype(np.float16), qs.astype(np.float16)
    proj = rng.standard_normal((d, dn)).astype(np.float32) / np.sqrt(d)
    nav = data @ proj; nav /= np.linalg.norm(nav, axis=1, keepdims=True)          # search.py:50-52
    qnav = qs @ proj; qnav /= np.linalg.norm(qnav, axis=1, keepdims=True)
    dims = {"MLP": [128], "MLP-5": [256, 128]}[model_type]
    w, b, prev = [], [], dn
    for h in dims + [C]:
        w.append((rng.standard_normal((h, prev)) / np.sqrt(prev)).astype(np.float32) * 2.0)
        b.append((0.1 * rng.standard_normal(h)).astype(np.float32))
        prev = h
    li = ref_shims.make_reference_index(w, b, model_type)
    ref = ref_shims.load_reference()
    data_nav = pd.DataFrame(nav); data_nav.index += 1                              # search.py:71-72
    data_search = pd.DataFrame(data16); data_search.index += 1                     # search.py:83-84
    pred_categories = li.model.predict(ref.model.data_X_to_torch(data_nav))        # LearnedIndex.py:240
    return dict(li=li, ref=ref, w=w, b=b, data_nav=data_nav, data_search=data_search, nav=nav, qnav=qnav,
                data16=data16, q16=q16, pc=np.asarray(pred_categories).astype(np.int64), C=C, nq=nq)


@pytest.mark.parametrize("seed,n,nq,C,model_type,nb", [(11, 3000, 150, 20, "MLP", 3), (12, 2500, 120, 14, "MLP-5", 4),
                                                      (13, 1200, 90, 30, "MLP", 6)])
def test_oracle_equals_live_reference(seed, n, nq, C, model_type, nb):
    c = _case(seed, n, nq, C, model_type)
    li, ref = c["li"], c["ref"]
    # ---- navigation (model.py:214-229, :201-212) ----
    probs, classes = li.model.predict_proba(ref.model.data_X_to_torch(c["qnav"]))
    o_probs, o_classes = O.predict_proba(c["qnav"], c["w"], c["b"])
    np.testing.assert_allclose(o_probs, probs, rtol=2e-5, atol=1e-7)
    assert (o_classes[:, :nb] == classes[:, :nb]).mean() > 0.995
    assert (O.predict(c["nav"], c["w"], c["b"]) == c["pc"]).mean() > 0.999
    classes = classes.astype(np.int64)
    sizes = np.bincount(c["pc"], minlength=C)
    small = ((sizes > 0) & (sizes < K))[classes[:, :nb]].any(axis=1)              # '<k' branch fires on the clean path
    # ---- search(use_threshold=False): the clean oracle (LearnedIndex.py:22-101) ----
    d, a = li.search(c["data_nav"].copy(), c["qnav"], c["data_search"], c["q16"], c["pc"], n_buckets=nb, k=K,
                     use_threshold=False)
    od, oa = O.search_intended(classes, c["data16"], c["q16"], c["pc"], nb, K)
    ok = ~small
    assert ok.sum() > 0
    assert_ids_equal_up_to_ties(oa[ok], od[ok], a[ok], d[ok], what="intended vs live reference")
    np.testing.assert_allclose(od[ok], d[ok], rtol=0, atol=1e-12)
    table = O.BucketTable(c["pc"], 1)
    ld = la = None
    for rank in range(nb):                                                         # the faithful loops, same merge
        dd, aa = O.search_single_loops(table, c["data16"], c["q16"], classes[:, rank], K, None)
        if la is None:
            ld, la = dd, aa
        else:
            dc, ac = np.hstack((ld, dd)), np.hstack((la, aa))
            sel = np.argsort(dc, axis=1, kind="stable")[:, :K]
            ld, la = np.take_along_axis(dc, sel, 1), np.take_along_axis(ac, sel, 1)
    assert_ids_equal_up_to_ties(la, ld, a, d, what="loops vs live reference")
    np.testing.assert_allclose(ld, d, rtol=0, atol=1e-12)
    # ---- search(use_threshold=True): what run() calls (search.py:119-128); D1 rows flagged ----
    with ref_shims.D1Flagger(nq, K) as fl:
        ds, as_ = li.search(c["data_nav"].copy(), c["qnav"], c["data_search"], c["q16"], c["pc"], n_buckets=nb, k=K,
                            use_threshold=True)
    ok2 = ok & ~fl.flags & (ds < O.EMPTY_DIST).all(axis=1)
    assert ok2.sum() > 0
    assert_ids_equal_up_to_ties(oa[ok2], od[ok2], as_[ok2], ds[ok2], what="intended vs shipped (unflagged rows)")
    # ---- search_single with a threshold (LearnedIndex.py:149-163, utils.py:14-43) ----
    nav_copy = c["data_nav"].copy(); nav_copy["category"] = c["pc"]
    d0, a0 = li.search_single(nav_copy, c["data_search"], c["q16"], classes[:, 0], k=K)
    thr = np.median(d0, axis=1)
    with ref_shims.D1Flagger(nq, K) as fl2:
        dt, at = li.search_single(nav_copy, c["data_search"], c["q16"], classes[:, 0], k=K, threshold_dist=thr)
    odt, oat = O.search_intended(classes[:, :1], c["data16"], c["q16"], c["pc"], 1, K, thresholds=thr)
    ok3 = ~fl2.flags
    live_ref, live_o = dt[ok3] < O.EMPTY_DIST, odt[ok3] < O.EMPTY_DIST
    assert (live_ref == live_o).all()
    assert_ids_equal_up_to_ties(np.where(live_o, oat[ok3], 0), odt[ok3], np.where(live_ref, at[ok3], 0), dt[ok3],
                                what="threshold path")
    np.testing.assert_allclose(odt[ok3], dt[ok3], rtol=0, atol=1e-12)
