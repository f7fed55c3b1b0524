"""Exactness as a guarantee: K3 selects candidates with fp32 scores and K5 issues a certificate per
query; when the certificate fails the query is re-derived from every probed object in float64 by
``exact_fixup_kernel`` instead of being returned on an empirical error bound.  Here near-ties are made denser than
the k' - k spare candidates on purpose (40 near-duplicates of every query, ``kprime_extra = 0``), so the certificate
must fail and the fallback must produce the oracle's rows."""
import numpy as np
import pytest

from conftest import assert_ids_equal_up_to_ties

pytestmark = pytest.mark.gpu


def _near_tie_case(seed=3, n=6000, nq=48, C=6, dup=40):
    rng = np.random.default_rng(seed)
    d, dn = 768, 96
    x = rng.standard_normal((n, d)).astype(np.float32)
    x /= np.linalg.norm(x, axis=1, keepdims=True)
    q = rng.standard_normal((nq, d)).astype(np.float32)
    q /= np.linalg.norm(q, axis=1, keepdims=True)
    pc = rng.integers(0, C, n).astype(np.int64)
    qcat = rng.integers(0, C, nq)
    for i in range(nq):                                  # 40 near-duplicates of query i inside its first bucket
        rows = np.where(pc == qcat[i])[0][i * dup % 500: i * dup % 500 + dup]
        x[rows] = q[i] + 3e-3 * rng.standard_normal((rows.shape[0], d)).astype(np.float32) / np.sqrt(d)
    x[10] = x[11]                                        # exact duplicates: ties resolve to the lower id
    qn = rng.standard_normal((nq, dn)).astype(np.float32)
    # a 1-layer "MLP" whose top class is qcat[i]: logits = W qn, solved per query by construction
    w = np.zeros((C, dn), np.float32); b = np.zeros(C, np.float32)
    qn = np.zeros((nq, dn), np.float32)
    for i in range(nq):
        qn[i, qcat[i]] = 5.0; qn[i, (qcat[i] + 1) % C] = 3.0; qn[i, (qcat[i] + 2) % C] = 1.0
    w[np.arange(C), np.arange(C)] = 1.0
    return dict(x=x.astype(np.float16), q=q.astype(np.float16), qn=qn, pc=pc, w=[w], b=[b], C=C, qcat=qcat)


@pytest.mark.parametrize("rescore", [3, 1, 0], ids=["tcgen05-pair", "tcgen05-tmemA", "simt"])
def test_uncertified_queries_are_recomputed_exactly(rescore):
    from lmi_b200.index import GpuLearnedIndex
    from oracle import lmi_oracle as O
    c = _near_tie_case()
    li = GpuLearnedIndex(c["C"], rescore=rescore).set_model(c["w"], c["b"])
    li.set_option("kprime_extra", 0)                      # k' = 16 for k = 10: only 6 spare candidates
    li.prepare(c["x"], c["pc"])
    _, classes = O.predict_proba(c["qn"], c["w"], c["b"])
    assert (classes[:, 0] == c["qcat"]).all()
    for nb, k in ((1, 10), (3, 10), (2, 16)):
        d, a = li._search_prepared(c["qn"], c["q"], nb, k)
        st = li.stats()
        assert st["n_uncertified"] >= c["q"].shape[0] // 2, st      # the certificate did fail for the near-tie queries
        od, oa = O.search_intended(classes, c["x"], c["q"], c["pc"], nb, k)
        assert_ids_equal_up_to_ties(a, d, oa, od, what=f"fallback nb={nb} k={k}")
        np.testing.assert_allclose(d, od, rtol=1e-3, atol=1e-9)
        assert (np.diff(d, axis=1) >= 0).all()
    # search_single with a threshold goes through the same fallback (strict '<', utils.py:23); the threshold sits between
    # the row's 6th and 7th distance, so exactly six slots survive and they are the oracle's
    d1, a1 = li.search_single(None, None, c["q"], c["qcat"], k=10)
    thr = 0.5 * (d1[:, 5] + d1[:, 6])
    sep = (d1[:, 6] - d1[:, 5]) > 1e-12
    dt, at = li.search_single(None, None, c["q"], c["qcat"], k=10, threshold_dist=thr)
    assert li.stats()["n_uncertified"] > 0
    odt, oat = O.search_intended(c["qcat"][:, None], c["x"], c["q"], c["pc"], 1, 10, thresholds=thr)
    assert_ids_equal_up_to_ties(at[sep], dt[sep], oat[sep], odt[sep], what="fallback with threshold")
    assert ((dt[sep] < 10000.0).sum(axis=1) == 6).all() and (at[sep][:, 6:] == 0).all()


def test_fallback_in_sharded_partials_matches_single():
    """The exchange entries a rank emits for an uncertified query come from the fallback too, and the merged result
    is bit-identical to the unsharded one (same float64 arithmetic on both paths)."""
    import torch
    from lmi_b200.index import GpuLearnedIndex
    c = _near_tie_case(seed=5)
    nb, k = 3, 10
    ref = GpuLearnedIndex(c["C"]).set_model(c["w"], c["b"]); ref.set_option("kprime_extra", 0); ref.prepare(c["x"], c["pc"])
    rd, ra = ref._search_prepared(c["qn"], c["q"], nb, k)
    assert ref.stats()["n_uncertified"] > 0
    # the same query set answered WITHOUT forcing failures (k' = 16 + default spare): the certified path and the fallback agree
    cert = GpuLearnedIndex(c["C"]).set_model(c["w"], c["b"]); cert.set_option("kprime_extra", 40); cert.prepare(c["x"], c["pc"])
    cd, ca = cert._search_prepared(c["qn"], c["q"], nb, k)
    assert (ca == ra).all() and (cd == rd).all()
    nq = c["q"].shape[0]
    qn, qs = torch.from_numpy(c["qn"]).cuda(), torch.from_numpy(c["q"]).cuda()
    for world, policy in ((2, 0), (3, 1)):
        parts, keep = [], []
        for r in range(world):
            li = GpuLearnedIndex(c["C"], rank=r, world=world, shard_policy=policy).set_model(c["w"], c["b"])
            li.set_option("kprime_extra", 0)
            li.prepare(c["x"], c["pc"])
            p = torch.empty(li.partial_bytes(nq, k), dtype=torch.uint8, device="cuda")
            li.search_partial_device(qn, qs, nb, k, p)
            parts.append(p); keep.append(li)
        out = (torch.empty(nq, k, dtype=torch.float64, device="cuda"), torch.empty(nq, k, dtype=torch.int32, device="cuda"))
        keep[0].merge_partials_device(torch.cat(parts), world, nq, k, out)
        torch.cuda.synchronize()
        assert (out[1].cpu().numpy().view(np.uint32) == ra).all()
        assert (out[0].cpu().numpy() == rd).all()
