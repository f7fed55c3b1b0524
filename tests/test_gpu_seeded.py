"""GPU parity on seeded synthetic inputs at sizes the oracle finishes in seconds, the edge cases
the reference special-cases or mishandles, the two rescoring kernels
against each other, and shard-merge equality (multi-GPU logic emulated on one device)."""
import numpy as np
import pytest

from conftest import assert_ids_equal_up_to_ties, TS_KERNELS

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def case():
    """40k x 768 LAION-shaped objects, 48 categories, MLP-5 trained with the fast trainer."""
    import torch
    from lmi_b200.synth import SynthSpec, make_synthetic
    from lmi_b200.trainer import train_index
    from oracle import lmi_oracle as O
    spec = SynthSpec(n=40_000, nq=600, n_blobs=400, sigma=0.6, pca_rows=20_000)
    t = make_synthetic(spec, "cuda")
    _, w, b = train_index(t["data_nav"], 48, "MLP-5", sample=40_000, epochs=4)
    a = {k: v.cpu().numpy() for k, v in t.items()}
    pc = O.predict(a["data_nav"], w, b)
    _, classes = O.predict_proba(a["queries_nav"], w, b)
    return dict(a, weights=w, biases=b, pred_categories=pc, classes=classes, C=48)


def _index(c, rescore=1, **kw):
    from lmi_b200.index import GpuLearnedIndex
    li = GpuLearnedIndex(c["C"], rescore=rescore, **kw)
    li.set_model(c["weights"], c["biases"])
    li.prepare(c["data_f16"], c["pred_categories"])
    return li


def _search(li, c, nb, k, sl=slice(None)):
    return li.search(None, c["queries_nav"][sl], c["data_f16"], c["queries_f16"][sl], c["pred_categories"],
                     n_buckets=nb, k=k)


def _rows_with_same_buckets(li, c, nb):
    """rows whose GPU bucket set equals the oracle's (they differ only on fp32 near-ties)."""
    _, cls = li.model.predict_proba(c["queries_nav"], n_top=nb)
    same = (np.sort(cls, 1) == np.sort(c["classes"][:, :nb], 1)).all(1)
    assert same.mean() > 0.99
    return same


@pytest.mark.parametrize("nb,k", [(1, 10), (4, 10), (7, 10), (4, 1), (3, 50), (2, 100)])
def test_seeded_vs_oracle(case, nb, k):
    from oracle import lmi_oracle as O
    c = case
    li = _index(c, 1)
    d, a = _search(li, c, nb, k)
    ok = _rows_with_same_buckets(li, c, nb)
    od, oa = O.search_intended(c["classes"], c["data_f16"], c["queries_f16"], c["pred_categories"], nb, k)
    assert_ids_equal_up_to_ties(a[ok], d[ok], oa[ok], od[ok], what=f"seeded nb={nb} k={k}")
    np.testing.assert_allclose(d[ok], od[ok], rtol=1e-3, atol=1e-9)
    st = li.stats()
    assert st["n_uncertified"] == 0 and st["pairs"] > 0, st
    # size-independent properties
    assert (np.diff(d, axis=1) >= 0).all()                                  # ascending
    for row_ids in a:
        nz = row_ids[row_ids != 0]
        assert len(set(nz.tolist())) == len(nz)                             # no duplicates (D1 would produce them)
    d2, a2 = _search(li, c, nb, k)
    assert (a2 == a).all() and (d2 == d).all()                              # idempotent / deterministic


@pytest.mark.parametrize("k", [10, 50])
def test_search_rank_threshold_vs_oracle(case, k):
    """lmi_search_rank with per-query thresholds (the reference's search_single(threshold_dist=...),
    LearnedIndex.py:149-163 / utils.py:22-23) against the oracle's thresholded semantics at 40k objects:
    the running threshold search() would use for rank 1, and a per-query random quantile of rank 0's distances."""
    from oracle import lmi_oracle as O
    c = case
    li = _index(c)                                  # library default kernel
    d0, a0 = li.search_single(None, c["data_f16"], c["queries_f16"], c["classes"][:, 0], k=k)
    rng = np.random.default_rng(5)
    # thresholds lie BETWEEN two distances of the row (a threshold that equals a candidate's distance to the last bit
    # would fall on either side of the strict '<' depending on float64 summation order)
    j = rng.integers(1, k, d0.shape[0])
    mid = 0.5 * (d0[np.arange(d0.shape[0]), j - 1] + d0[np.arange(d0.shape[0]), j])
    sep = (d0[np.arange(d0.shape[0]), j] - d0[np.arange(d0.shape[0]), j - 1]) > 1e-9
    for cats, thr, ok in ((c["classes"][:, 1], d0.max(axis=1), np.ones(d0.shape[0], bool)), (c["classes"][:, 0], mid, sep)):
        d, a = li.search_single(None, c["data_f16"], c["queries_f16"], cats, k=k, threshold_dist=thr)
        od, oa = O.search_intended(cats[:, None], c["data_f16"], c["queries_f16"], c["pred_categories"], 1, k, thresholds=thr)
        ok = ok & ~(np.abs(od - thr[:, None]) < 1e-12).any(axis=1)
        assert ok.mean() > 0.95
        assert_ids_equal_up_to_ties(a[ok], d[ok], oa[ok], od[ok], what=f"search_rank threshold k={k}")
        np.testing.assert_allclose(d[ok], od[ok], rtol=1e-3, atol=1e-9)
        assert (a[d == 10000.0] == 0).all()
        assert li.stats()["n_uncertified"] == 0
    # second case: exactly j survivors per row (everything below the midpoint)
    assert ((d[ok] < 10000.0).sum(axis=1) == j[ok]).all()


@pytest.mark.parametrize("nb,k", [(4, 10), (3, 50), (2, 100)])
def test_tc_equals_simt(case, nb, k):
    c = case
    li1 = _index(c, 1)
    d1, a1 = _search(li1, c, nb, k)
    assert li1.stats()["rescore_kernel"] in TS_KERNELS      # k = 100 (k' = 128) stays on tcgen05 too
    d0, a0 = _search(_index(c, 0), c, nb, k)
    assert (a1 == a0).all()
    assert (d1 == d0).all()            # same float64 refine on the same survivors -> bit-identical


def test_recall_vs_brute_force(case):
    from oracle import lmi_oracle as O
    c = case
    sub = slice(0, 200)
    gt_d, gt_i = O.brute_force(c["queries_f16"][sub], c["data_f16"], 10)
    li = _index(c, 1)
    r = []
    for nb in (1, 4, 16, 48):
        _, a = _search(li, c, nb, 10, sub)
        r.append(O.get_recall(a, gt_i, 10))
    assert r[0] <= r[1] <= r[2] <= r[3]
    assert r[3] == 1.0                                                       # exhaustive probe = brute force


def test_edge_cases():
    """empty category at some rank, bucket with < k objects, category with exactly one query,
    zero-norm vectors, n_buckets=1, n_buckets > #non-empty buckets, ragged nq."""
    from oracle import lmi_oracle as O
    from lmi_b200.index import GpuLearnedIndex
    rng = np.random.default_rng(7)
    n, C, nq, d, dn = 3000, 12, 37, 768, 96
    data = rng.standard_normal((n, d)).astype(np.float16)
    data[5] = 0                                                  # zero-norm object -> cos 0, dist 1
    q = rng.standard_normal((nq, d)).astype(np.float16)
    q[3] = 0                                                     # zero-norm query
    qn = rng.standard_normal((nq, dn)).astype(np.float32)
    w = [rng.standard_normal((20, dn)).astype(np.float32) * 0.3, rng.standard_normal((C, 20)).astype(np.float32)]
    b = [np.zeros(20, np.float32), np.zeros(C, np.float32)]
    pc = rng.integers(0, C, n)
    pc[pc == 4] = 5                                              # category 4 empty
    pc[pc == 7] = 8
    pc[:6] = 7                                                   # category 7 holds 6 (< k) objects, incl. the zero vector
    pc[pc == 9] = 10
    pc[100] = 9                                                  # category 9 holds exactly one object
    _, classes = O.predict_proba(qn, w, b)
    for rescore in (0, 1, 2):
        li = GpuLearnedIndex(C, rescore=rescore).set_model(w, b)
        for nb in (1, 3, C):
            d_, a_ = li.search(None, qn, data, q, pc, n_buckets=nb, k=10)
            od, oa = O.search_intended(classes, data, q, pc, nb, 10)
            assert_ids_equal_up_to_ties(a_, d_, oa, od, tie_eps=1e-9, what=f"edge nb={nb} rescore={rescore}")
            np.testing.assert_allclose(d_, od, rtol=1e-3, atol=1e-9)
        # zero-norm query: every distance is exactly 1 (scikit-learn maps a zero norm to scale 1)
        assert np.allclose(d_[3], 1.0)
        # single rank into the tiny / empty buckets through search_single
        cat = np.full(nq, 7); cat[0] = 4; cat[1] = 9
        ds, as_ = li.search_single(None, data, q, cat, k=10)
        assert (as_[0] == 0).all() and (ds[0] == 10000.0).all()                 # empty category -> nothing
        assert (as_[1, 0] == 101) and (as_[1, 1:] == 0).all()                   # one object (1-based id)
        assert (as_[2, :6] != 0).all() and (as_[2, 6:] == 0).all() and (ds[2, 6:] == 10000.0).all()
        assert set(as_[2, :6].tolist()) == {1, 2, 3, 4, 5, 6}


@pytest.mark.parametrize("nb,k", [(4, 10), (2, 100)], ids=["k10", "k100"])
@pytest.mark.parametrize("policy", [0, 1], ids=["buckets", "stripes"])
@pytest.mark.parametrize("world", [2, 3, 8])
def test_sharded_merge_equals_single(case, world, policy, nb, k):
    """Every rank holds a shard; per-rank partial results merged by the merge kernel must equal the
    one-GPU result bit for bit (ids and distances).  Ranks are emulated as `world` indexes on one
    device; the all-gather is a concatenation."""
    import torch
    from lmi_b200.index import GpuLearnedIndex
    c = case
    ref_d, ref_a = _search(_index(c, 1), c, nb, k)
    nq = c["queries_f16"].shape[0]
    qn = torch.from_numpy(c["queries_nav"]).cuda()
    qs = torch.from_numpy(c["queries_f16"]).cuda()
    parts, total = [], 0
    keep = []
    for r in range(world):
        li = GpuLearnedIndex(c["C"], rank=r, world=world, shard_policy=policy).set_model(c["weights"], c["biases"])
        li.prepare(c["data_f16"], c["pred_categories"])
        total += int(li.bucket_ptr()[-1])
        p = torch.empty(li.partial_bytes(nq, k), dtype=torch.uint8, device="cuda")
        li.search_partial_device(qn, qs, nb, k, p)
        parts.append(p); keep.append(li)
    assert total == c["data_f16"].shape[0]                       # shards partition the objects
    gathered = torch.cat(parts)
    out = (torch.empty(nq, k, dtype=torch.float64, device="cuda"), torch.empty(nq, k, dtype=torch.int32, device="cuda"))
    keep[0].merge_partials_device(gathered, world, nq, k, out)
    torch.cuda.synchronize()
    assert (out[1].cpu().numpy().view(np.uint32) == ref_a).all()
    assert (out[0].cpu().numpy() == ref_d).all()


@pytest.mark.parametrize("world", [2, 3])
def test_sliced_navigation_equals_replicated(case, world):
    """Multi-GPU navigation split (DistributedSearch(slice_navigation=True)): every rank navigates rows
    [r*per, (r+1)*per) with lmi_navigate_device, the int32 bucket lists are concatenated (the all-gather) and fed to
    lmi_search_partial_classes_device.  Lists and partial results must equal what full navigation on the rank gives."""
    import torch
    from lmi_b200.index import GpuLearnedIndex
    c = case
    nb, k = 4, 10
    nq = c["queries_f16"].shape[0]
    qn = torch.from_numpy(c["queries_nav"]).cuda()
    qs = torch.from_numpy(c["queries_f16"]).cuda()
    per = (nq + world - 1) // world
    li = GpuLearnedIndex(c["C"], rank=1, world=world).set_model(c["weights"], c["biases"])
    li.prepare(c["data_f16"], c["pred_categories"])
    every = torch.full((world * per, nb), -1, dtype=torch.int32, device="cuda")
    for r in range(world):
        lo, hi = min(nq, r * per), min(nq, (r + 1) * per)
        if hi > lo:
            li.navigate_device(qn[lo:hi], nb, every[r * per:r * per + (hi - lo)])
    _, full = li.model.predict_proba(c["queries_nav"], n_top=nb)
    assert (every[:nq].cpu().numpy() == full).all()
    p_ref = torch.empty(li.partial_bytes(nq, k), dtype=torch.uint8, device="cuda")
    li.search_partial_device(qn, qs, nb, k, p_ref)
    p_new = torch.empty_like(p_ref)
    li.search_partial_classes_device(every, qs, nb, k, p_new)
    torch.cuda.synchronize()
    assert torch.equal(p_new, p_ref)
    # ids outside [0, C) match no bucket: a list of -1 gives an empty row
    every[:8] = -1
    li.search_partial_classes_device(every, qs, nb, k, p_new)
    torch.cuda.synchronize()
    from lmi_b200.shard import PARTIAL_DTYPE
    part = np.frombuffer(p_new.cpu().numpy().tobytes(), dtype=PARTIAL_DTYPE).reshape(nq, k)
    assert (part["id"][:8] == 0).all() and (part["dist"][:8] == 10000.0).all()


@pytest.mark.parametrize("world", [2, 5])
def test_presharded_ranges_equal_single(case, world):
    """SHARD_RANGES: every rank loads only its own row range into an ordinary index (id_base = first
    row + 1) -- the 100M layout, where no process ever holds the whole dataset.  Merged partials must
    equal the one-GPU result bit for bit."""
    import torch
    from lmi_b200.index import GpuLearnedIndex
    from lmi_b200.shard import range_of_rank
    c = case
    nb, k = 4, 10
    ref_d, ref_a = _search(_index(c, 1), c, nb, k)
    nq, n = c["queries_f16"].shape[0], c["data_f16"].shape[0]
    qn = torch.from_numpy(c["queries_nav"]).cuda()
    qs = torch.from_numpy(c["queries_f16"]).cuda()
    x = torch.from_numpy(c["data_f16"]).cuda()
    pc = torch.from_numpy(c["pred_categories"].astype(np.int64)).cuda()
    parts, keep = [], []
    for r in range(world):
        lo, hi = range_of_rank(n, r, world)
        li = GpuLearnedIndex(c["C"]).set_model(c["weights"], c["biases"])
        li.prepare_device(x[lo:hi].contiguous(), pc[lo:hi].contiguous(), id_base=lo + 1)
        p = torch.empty(li.partial_bytes(nq, k), dtype=torch.uint8, device="cuda")
        li.search_partial_device(qn, qs, nb, k, p)
        parts.append(p); keep.append(li)
    out = (torch.empty(nq, k, dtype=torch.float64, device="cuda"), torch.empty(nq, k, dtype=torch.int32, device="cuda"))
    keep[0].merge_partials_device(torch.cat(parts), world, nq, k, out)
    torch.cuda.synchronize()
    assert (out[1].cpu().numpy().view(np.uint32) == ref_a).all()
    assert (out[0].cpu().numpy() == ref_d).all()


@pytest.mark.parametrize("rescore", [0, 1, 2], ids=["simt", "tcgen05-tmemA", "tcgen05-ss"])
def test_small_chunks_many_work_items(case, rescore):
    """chunk_objects=512 splits every bucket into several work items and partial lists; the result
    must not depend on the chunking."""
    c = case
    ref_d, ref_a = _search(_index(c, 1), c, 5, 10)
    d, a = _search(_index(c, rescore, chunk_objects=512), c, 5, 10)
    assert (a == ref_a).all() and (d == ref_d).all()


def test_config0_300k_shape():
    """BASELINE.json configs[0] at its real shape -- 300K objects, C = 122, `-bp 6` -> 7 leaf buckets
    (search/search.py:37-38), k = 10 -- on a 1,000-query sample the oracle finishes in seconds: ids exact
    up to documented ties, distances to 1e-3 relative (observed ~1e-15), every query certified."""
    import torch
    from lmi_b200.synth import SynthSpec, make_synthetic
    from lmi_b200.trainer import train_index
    from lmi_b200.index import GpuLearnedIndex
    from oracle import lmi_oracle as O
    C, nb, k = 122, int(6 / 100 * 122), 10
    assert nb == 7
    spec = SynthSpec(n=300_000, nq=1_000, n_blobs=2000, sigma=0.55)
    t = make_synthetic(spec, "cuda")
    _, w, b = train_index(t["data_nav"], C, "MLP-5", sample=200_000, epochs=3)
    a = {kk: v.cpu().numpy() for kk, v in t.items()}
    pc = O.predict(a["data_nav"], w, b)
    _, classes = O.predict_proba(a["queries_nav"], w, b)
    li = GpuLearnedIndex(C).set_model(w, b)
    assert (li.model.predict(a["data_nav"]) != pc).mean() < 1e-4          # N1: argmax bucket of all 300K objects (fp32 near-ties aside)
    li.prepare(a["data_f16"], pc)
    d, ids = li.search(None, a["queries_nav"], a["data_f16"], a["queries_f16"], pc, n_buckets=nb, k=k)
    _, cls = li.model.predict_proba(a["queries_nav"], n_top=nb)
    ok = (np.sort(cls, 1) == np.sort(classes[:, :nb], 1)).all(1)
    assert ok.mean() > 0.99
    od, oa = O.search_intended(classes, a["data_f16"], a["queries_f16"], pc, nb, k)
    assert_ids_equal_up_to_ties(ids[ok], d[ok], oa[ok], od[ok], what="config 0 (300K, 7 buckets)")
    np.testing.assert_allclose(d[ok], od[ok], rtol=1e-3, atol=1e-9)
    st = li.stats()
    assert st["n_uncertified"] == 0 and st["rescore_kernel"] in ("rescore_ts_kernel", "rescore_ts2_kernel")   # library default (unsharded: pair kernel)


@pytest.mark.parametrize("nb,k", [(1, 10), (2, 100)])
def test_short_items_repeated_launches(case, nb, k):
    """Many back-to-back launches over short work items (one query tile per bucket -> plain pairs and
    dummy entries, 256-object chunks): the kernel's cross-CTA barrier protocol must neither fault nor
    change the answer.  (Round 1 had a rare fault here: an MMA warp waited for a stage phase it had not
    seen the predecessor of, and CTAs could exit with the peer's commit arrives still in flight.)"""
    import torch
    from lmi_b200.index import GpuLearnedIndex
    c = case
    li = GpuLearnedIndex(c["C"], rank=3, world=8, chunk_objects=256).set_model(c["weights"], c["biases"])
    li.prepare(c["data_f16"], c["pred_categories"])
    qn = torch.from_numpy(c["queries_nav"]).cuda()
    qs = torch.from_numpy(c["queries_f16"]).cuda()
    nq = qs.shape[0]
    first = torch.empty(li.partial_bytes(nq, k), dtype=torch.uint8, device="cuda")
    li.search_partial_device(qn, qs, nb, k, first)
    p = torch.empty_like(first)
    for _ in range(300):
        li.search_partial_device(qn, qs, nb, k, p)
    torch.cuda.synchronize()
    assert torch.equal(p, first)
    assert li.stats()["n_uncertified"] == 0
