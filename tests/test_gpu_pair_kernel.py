"""GPU parity of the CTA-pair rescoring kernel (option rescore = 3, ``rescore_ts2_kernel``: tcgen05
``cta_group::2`` MMAs, planner pair mode) against the default kernel, the oracle and the reference's
golden vectors.  Every rescoring kernel hands the same survivors to the same float64 refine, so the
final (dists, ids) must be BIT-IDENTICAL to the default kernel's.

The kernel is opt-in this round ("K3-pair").  ``LMI_FORCE_PAIR_KERNEL=1`` additionally
reroutes every default-kernel caller of the whole GPU suite to it (lmi_b200.index).
"""
import numpy as np
import pytest

from conftest import load_golden, GOLDEN_TAGS, rows_with_small_buckets, assert_ids_equal_up_to_ties

pytestmark = pytest.mark.gpu

PAIR = 3


def _golden_index(g, rescore, **kw):
    from lmi_b200.index import GpuLearnedIndex
    li = GpuLearnedIndex(g["n_categories"], d_nav=g["queries_nav"].shape[1], d=g["data_f16"].shape[1],
                         rescore=rescore, **kw)
    li.set_model(g["weights"], g["biases"])
    return li


@pytest.mark.parametrize("tag", GOLDEN_TAGS)
def test_pair_kernel_matches_reference_and_oracle(tag):
    from oracle import lmi_oracle as O
    g = load_golden(tag)
    li = _golden_index(g, PAIR)
    for nb in g["bucket_list"]:
        d, a = li.search(None, g["queries_nav"], g["data_f16"], g["queries_f16"], g["pred_categories"],
                         n_buckets=nb, k=g["k"])
        st = li.stats()
        assert st["rescore_kernel"] == "rescore_ts2_kernel", st
        ok = ~rows_with_small_buckets(g, nb)
        assert_ids_equal_up_to_ties(a[ok], d[ok], g[f"clean_ids_nb{nb}"][ok], g[f"clean_dists_nb{nb}"][ok],
                                    what=f"{tag} nb={nb} pair kernel vs reference")
        od, oa = O.search_intended(g["classes"], g["data_f16"], g["queries_f16"], g["pred_categories"], nb, g["k"])
        assert_ids_equal_up_to_ties(a, d, oa, od, what=f"{tag} nb={nb} pair kernel vs oracle")
        np.testing.assert_allclose(d, od, rtol=1e-3, atol=1e-9)
        assert st["n_uncertified"] == 0 and st["max_abs_score_err"] < 5e-6, st


@pytest.fixture(scope="module")
def case():
    """40k x 768 LAION-shaped objects, 48 categories, MLP-5 (same recipe as test_gpu_seeded.case)."""
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


def _index(c, rescore, **kw):
    from lmi_b200.index import GpuLearnedIndex
    li = GpuLearnedIndex(c["C"], rescore=rescore, **kw)
    li.set_model(c["weights"], c["biases"])
    li.prepare(c["data_f16"], c["pred_categories"])
    return li


def _search(li, c, nb, k):
    return li.search(None, c["queries_nav"], c["data_f16"], c["queries_f16"], c["pred_categories"], n_buckets=nb, k=k)


@pytest.mark.parametrize("sets", [1, 2], ids=["1set", "2sets"])
@pytest.mark.parametrize("nb,k,chunk", [(1, 10, 0), (4, 10, 0), (7, 10, 0), (4, 10, 512), (48, 10, 256), (4, 20, 320),
                                        (3, 50, 0), (2, 100, 0), (2, 100, 1024)])
def test_pair_kernel_equals_default_kernel(case, nb, k, chunk, sets):
    """`sets`: epilogue warp sets of the pair kernel (option "epi_sets"; two sets exist for k' <= 32, larger k'
    always run one), each with its own per-row candidate list merged in shared memory at the end of an item."""
    c = case
    if sets == 2 and k + 6 > 32:
        pytest.skip("two epilogue sets exist for k' <= 32 only")
    li3 = _index(c, PAIR, chunk_objects=chunk)
    li3.set_option("epi_sets", sets)
    d3, a3 = _search(li3, c, nb, k)
    st = li3.stats()
    assert st["rescore_kernel"] == "rescore_ts2_kernel" and st["n_uncertified"] == 0, st
    assert st["n_work_items"] % 2 == 0
    d1, a1 = _search(_index(c, 1, chunk_objects=chunk), c, nb, k)
    assert (a3 == a1).all()
    assert (d3 == d1).all()
    # repeated launches: deterministic, and the persistent state (barriers, tensor memory) is released cleanly
    for _ in range(3):
        d_, a_ = _search(li3, c, nb, k)
        assert (a_ == a3).all() and (d_ == d3).all()


@pytest.mark.parametrize("order,hint", [(1, 0), (2, 0), (0, 1), (2, 1), (0, 2)],
                         ids=["bucket_major", "three_halves", "l2_hint", "three_halves_l2_hint", "lockstep"])
@pytest.mark.parametrize("nb,k,chunk", [(4, 10, 0), (24, 10, 256), (48, 10, 512)])   # 600 queries x nb / 48 buckets: ~1, ~3 and ~5 query tiles per bucket
def test_pair_kernel_work_list_variants(case, nb, k, chunk, order, hint):
    """Work-list orders of the CTA-pair kernel (option "item_order": 0 full pairs first -- the default --, 1 bucket-major,
    2 three-tile buckets as three half pairs) the L2 eviction hint on the chunk loads (option "l2_hint") and lock-step followers (option "lockstep") change when
    and how a chunk is streamed, never what is scored: ids and distances equal the default's bit for bit."""
    c = case
    ref = _index(c, PAIR, chunk_objects=chunk)
    d0, a0 = _search(ref, c, nb, k)
    li = _index(c, PAIR, chunk_objects=chunk)
    li.set_option("item_order", order)
    li.set_option("l2_hint", hint & 1)
    li.set_option("lockstep", hint >> 1)            # half pairs follow their chunk's full pair (bucket-major list)
    for _ in range(2):
        d, a = _search(li, c, nb, k)
        assert li.stats()["rescore_kernel"] == "rescore_ts2_kernel"
        assert (a == a0).all() and (d == d0).all()
    if order == 2:
        assert li.stats()["n_work_items"] >= ref.stats()["n_work_items"]


def test_pair_kernel_search_single(case):
    c = case
    li3, li1 = _index(c, PAIR), _index(c, 1)
    cat = c["classes"][:, 0]
    d3, a3 = li3.search_single(None, c["data_f16"], c["queries_f16"], cat, k=10)
    d1, a1 = li1.search_single(None, c["data_f16"], c["queries_f16"], cat, k=10)
    assert (a3 == a1).all() and (d3 == d1).all()
