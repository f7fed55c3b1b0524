"""GPU parity tests: the CUDA path, called through the C-ABI (lmi_b200.index -> liblmi_b200.so),
against (a) arrays returned by the reference itself (tests/golden/*.npz) and (b) the CPU oracle
on seeded inputs.  Bars: ids bit-exact except documented equal-distance ties (oracle distances
within 1e-12); distances within 1e-3 relative (BASELINE.json north_star) -- in practice ~1e-15.
"""
import numpy as np
import pytest

from conftest import load_golden, GOLDEN_TAGS, rows_with_small_buckets, assert_ids_equal_up_to_ties, TS_KERNELS

pytestmark = pytest.mark.gpu

DIST_RTOL = 1e-3      # tolerance stated by north_star
DIST_ATOL = 1e-9      # distances of near-duplicates are ~1e-16 in the reference; relative alone is meaningless there


def _index(g, rescore, **kw):
    from lmi_b200.index import GpuLearnedIndex
    li = GpuLearnedIndex(g["n_categories"], d_nav=g["queries_nav"].shape[1], d=g["data_f16"].shape[1],
                         rescore=rescore, **kw)
    li.set_model(g["weights"], g["biases"])
    return li


def _check_dists(d, ref):
    np.testing.assert_allclose(d, ref, rtol=DIST_RTOL, atol=DIST_ATOL)


@pytest.mark.parametrize("tag", GOLDEN_TAGS)
def test_navigation_matches_reference(tag):
    g = load_golden(tag)
    li = _index(g, 1)
    probs, classes = li.model.predict_proba(g["queries_nav"])
    assert probs.dtype == np.float32 and classes.dtype == np.int64 and classes.shape == g["classes"].shape
    np.testing.assert_allclose(probs, g["probs"], rtol=1e-4, atol=1e-7)
    nb = max(g["bucket_list"])
    # bucket order on the ranks the search consumes: identical unless two reference probabilities
    # are within fp32 rounding of each other (documented exception iii)
    diff = classes[:, :nb] != g["classes"][:, :nb]
    for r, c in zip(*np.where(diff)):
        p = g["probs"][r]
        assert abs(p[c] - p[min(c + 1, len(p) - 1)]) <= 1e-6 * p[c] or abs(p[c] - p[max(c - 1, 0)]) <= 1e-6 * p[c]
    assert diff.mean() < 0.01
    # every row is a permutation of all categories
    assert (np.sort(classes, axis=1) == np.arange(g["n_categories"])).all()
    # predict == argmax
    from oracle import lmi_oracle as O
    pc = li.model.predict(g["queries_nav"])
    ref = O.predict(g["queries_nav"], g["weights"], g["biases"])
    assert (pc == ref).mean() > 0.995


@pytest.mark.parametrize("rescore", [0, 1, 2], ids=["simt", "tcgen05-tmemA", "tcgen05-ss"])
@pytest.mark.parametrize("tag", GOLDEN_TAGS)
def test_search_matches_reference_clean(tag, rescore):
    """vs the reference's own search(..., use_threshold=False) on rows where it is defect-free,
    and vs the oracle's intended semantics on the rest."""
    from oracle import lmi_oracle as O
    g = load_golden(tag)
    li = _index(g, rescore)
    for nb in g["bucket_list"]:
        d, a = li.search(None, g["queries_nav"], g["data_f16"], g["queries_f16"], g["pred_categories"],
                         n_buckets=nb, k=g["k"])
        assert d.dtype == np.float64 and a.dtype == np.uint32 and d.shape == (g["queries_f16"].shape[0], g["k"])
        ok = ~rows_with_small_buckets(g, nb)
        assert_ids_equal_up_to_ties(a[ok], d[ok], g[f"clean_ids_nb{nb}"][ok], g[f"clean_dists_nb{nb}"][ok],
                                    what=f"{tag} nb={nb} rescore={rescore}")
        _check_dists(d[ok], g[f"clean_dists_nb{nb}"][ok])
        od, oa = O.search_intended(g["classes"], g["data_f16"], g["queries_f16"], g["pred_categories"], nb, g["k"])
        assert_ids_equal_up_to_ties(a, d, oa, od, what=f"{tag} nb={nb} vs oracle")
        _check_dists(d, od)
        st = li.stats()
        assert st["n_uncertified"] == 0, st
        assert st["max_abs_score_err"] < 5e-6, st
        assert st["rescore_kernel"] == ["rescore_simt_kernel", TS_KERNELS[0], "rescore_tc_kernel"][rescore]


@pytest.mark.parametrize("tag", GOLDEN_TAGS)
def test_search_matches_reference_shipped(tag):
    """vs what run() really calls (use_threshold=True, search.py:119-128): bit-exact ids on every
    row the reference's '<k' padding defect (D1) did not touch; touched rows are counted."""
    g = load_golden(tag)
    li = _index(g, 1)
    for nb in g["bucket_list"]:
        d, a = li.search(None, g["queries_nav"], g["data_f16"], g["queries_f16"], g["pred_categories"],
                         n_buckets=nb, k=g["k"], use_threshold=True)
        ok = ~g[f"shipped_d1_nb{nb}"] & ~rows_with_small_buckets(g, nb)
        ok &= (g[f"shipped_dists_nb{nb}"] < 10000.0).all(axis=1)
        assert ok.sum() > 0
        assert_ids_equal_up_to_ties(a[ok], d[ok], g[f"shipped_ids_nb{nb}"][ok], g[f"shipped_dists_nb{nb}"][ok],
                                    what=f"{tag} shipped nb={nb}")
        _check_dists(d[ok], g[f"shipped_dists_nb{nb}"][ok])


@pytest.mark.parametrize("rescore", [0, 1, 2], ids=["simt", "tcgen05-tmemA", "tcgen05-ss"])
@pytest.mark.parametrize("tag", GOLDEN_TAGS)
def test_search_single_matches_reference(tag, rescore):
    g = load_golden(tag)
    li = _index(g, rescore)
    li.prepare(g["data_f16"], g["pred_categories"])
    d, a = li.search_single(None, g["data_f16"], g["queries_f16"], g["classes"][:, 0], k=g["k"])
    ok = ~rows_with_small_buckets(g, 1)
    assert_ids_equal_up_to_ties(a[ok], d[ok], g["single_ids"][ok], g["single_dists"][ok], what=f"{tag} single")
    _check_dists(d[ok], g["single_dists"][ok])
    # threshold variant (utils.py:22-23: strict '<'): equals filtering the unthresholded result
    thr = np.median(d[:, : g["k"] // 2], axis=1)
    dt, at = li.search_single(None, g["data_f16"], g["queries_f16"], g["classes"][:, 0], k=g["k"], threshold_dist=thr)
    keep = d < thr[:, None]
    assert (at[keep] == a[keep]).all() and (dt[keep] == d[keep]).all()
    assert (at[~keep] == 0).all() and (dt[~keep] == 10000.0).all()


@pytest.mark.parametrize("rescore", [0, 1, 3], ids=["simt", "tcgen05-tmemA", "tcgen05-pair"])
@pytest.mark.parametrize("tag", GOLDEN_TAGS)
def test_search_single_threshold_matches_reference_and_oracle(tag, rescore):
    """Row a7 (utils.py:14-43 through LearnedIndex.py:149-163): ``lmi_search_rank(threshold)`` against (a) the oracle's
    intended semantics with the same thresholds -- every slot, pruned slots are (0, 10000.0) on both sides -- and
    (b) the arrays the reference's own ``search_single(threshold_dist=...)`` returned (tests/golden, made by
    oracle/make_golden.py), on the live slots (dist < 10000; the reference pads the others with arbitrary ids)
    of the rows its '<k' defect did not touch."""
    from oracle import lmi_oracle as O
    g = load_golden(tag)
    li = _index(g, rescore)
    li.prepare(g["data_f16"], g["pred_categories"])
    k = g["k"]
    for key, rank in (("single_thr", 1), ("single_thr2", 0)):
        thr, ref_d, ref_a, d1 = g[key], g[key + "_dists"], g[key + "_ids"], g[key + "_d1"]
        cats = g["classes"][:, rank]
        d, a = li.search_single(None, g["data_f16"], g["queries_f16"], cats, k=k, threshold_dist=thr)
        od, oa = O.search_intended(cats[:, None], g["data_f16"], g["queries_f16"], g["pred_categories"], 1, k, thresholds=thr)
        assert_ids_equal_up_to_ties(a, d, oa, od, what=f"{tag} {key} vs oracle")
        _check_dists(d, od)
        assert ((d < 10000.0) == (od < 10000.0)).all()
        assert (a[d == 10000.0] == 0).all()
        assert (d[d < 10000.0] < np.repeat(thr[:, None], k, 1)[d < 10000.0]).all()          # strict '<' (utils.py:23)
        ok = ~d1
        live = ref_d[ok] < 10000.0
        assert ((d[ok] < 10000.0) == live).all(), f"{tag} {key}: survivors per row differ from the reference"
        assert_ids_equal_up_to_ties(np.where(live, a[ok], 0), d[ok], np.where(live, ref_a[ok], 0), ref_d[ok],
                                    what=f"{tag} {key} vs reference")
        _check_dists(d[ok], ref_d[ok])
        assert li.stats()["n_uncertified"] == 0


def test_exhaustive_equals_reference_baseline():
    """n_buckets = C probes every bucket: must equal the reference's brute force (Baseline.py:14-20)."""
    g = load_golden("ref_mlp5_n1600_c16")
    li = _index(g, 1)
    sub = g["baseline_ids"].shape[0]
    d, a = li.search(None, g["queries_nav"][:sub], g["data_f16"], g["queries_f16"][:sub], g["pred_categories"],
                     n_buckets=g["n_categories"], k=g["k"])
    assert_ids_equal_up_to_ties(a, d, g["baseline_ids"], g["baseline_dists"], what="exhaustive")
    _check_dists(d, g["baseline_dists"])


def test_gpu_baseline_equals_reference_baseline():
    """N4: GpuBaseline.search mirrors li.Baseline.search (Baseline.py:14-20) -- same ids (1-based row
    numbers), float64 distances, and a duration as third return value."""
    from lmi_b200.index import GpuBaseline
    g = load_golden("ref_mlp5_n1600_c16")
    sub = g["baseline_ids"].shape[0]
    bl = GpuBaseline(device=0)
    d, a, secs = bl.search(g["queries_f16"][:sub], g["data_f16"], k=g["k"])
    assert secs > 0 and a.dtype == np.uint32 and d.dtype == np.float64
    assert_ids_equal_up_to_ties(a, d, g["baseline_ids"], g["baseline_dists"], what="GpuBaseline")
    _check_dists(d, g["baseline_dists"])
    d2, a2, _ = bl.search(g["queries_f16"][:sub], g["data_f16"], k=g["k"])        # store reused
    assert np.array_equal(a, a2) and np.array_equal(d, d2)


def test_errors_mirror_reference():
    g = load_golden("ref_mlp5_n1600_c16")
    from lmi_b200.index import GpuLearnedIndex
    li = GpuLearnedIndex(g["n_categories"])
    with pytest.raises(AssertionError):          # LearnedIndex.py:60
        li.search(None, g["queries_nav"], g["data_f16"], g["queries_f16"], g["pred_categories"], n_buckets=2)
    li.set_model(g["weights"], g["biases"])
    with pytest.raises(ValueError):
        li.search(None, g["queries_nav"], g["data_f16"], g["queries_f16"], g["pred_categories"], n_buckets=0)
    with pytest.raises(ValueError):
        li.prepare(g["data_f16"], g["pred_categories"][:-1])
    bad = g["pred_categories"].copy(); bad[3] = g["n_categories"] + 5
    with pytest.raises(ValueError):
        li.prepare(g["data_f16"], bad)
