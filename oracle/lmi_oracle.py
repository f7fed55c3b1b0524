"""CPU oracle for the LMI query-time hot path  --  TEST INFRASTRUCTURE, NOT PRODUCT CODE.

This file is a numpy restatement of the algorithm the reference implements in
``search/li/LearnedIndex.py``, ``search/li/model.py`` and ``search/li/utils.py``.
Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs may import it, and only as the checker / timed CPU arm.
The product path (``lmi_b200``) never imports anything from ``oracle/``.

Pinning: the reference ships no tests or golden vectors, so the
oracle is pinned against outputs of the *reference itself*, executed in the build
container through ``oracle/ref_shims.py`` and frozen by ``oracle/make_golden.py`` into
``tests/golden/*.npz``.  ``tests/test_oracle_golden.py`` replays every fixture.

Third-party arithmetic that the reference delegates to (none of it pinned by the
reference) is restated from its published algorithm:

* ``torch.nn.Linear`` / ``ReLU`` / ``softmax`` / ``topk``    -> ``mlp_logits``, ``predict_proba``
* ``sklearn.metrics.pairwise.cosine_similarity`` (upcast to float64, divide every row by
  its L2 norm with zero norms replaced by 1, then one matrix product)  -> ``cosine_distance``
* ``numpy.argsort`` (quicksort per bucket, stable for the rank merge)  -> used directly
"""
from __future__ import annotations

import numpy as np

EMPTY_DIST = 10_000.0  # reference: LearnedIndex.py:138 (fill value of never-written slots)


# --------------------------------------------------------------------------------------
# navigation (reference: search/li/model.py:15-83 Model, :86-89 data_X_to_torch,
#             :214-229 NeuralNetwork.predict_proba, :201-212 NeuralNetwork.predict)
# --------------------------------------------------------------------------------------
def mlp_logits(x, weights, biases):
    """Forward pass of ``Sequential(Linear, ReLU, ..., Linear)`` in float32.

    reference: model.py:21-78 (layer stacks), model.py:81-83 (forward).  ``weights[i]`` is
    the ``nn.Linear.weight`` of layer i, shape [out, in], ``biases[i]`` shape [out].
    """
    h = np.asarray(x).astype(np.float32)           # model.py:88
    n = len(weights)
    for i, (w, b) in enumerate(zip(weights, biases)):
        h = h @ np.asarray(w, dtype=np.float32).T + np.asarray(b, dtype=np.float32)
        if i + 1 < n:
            h = np.maximum(h, np.float32(0))        # torch.nn.ReLU
    return h


def softmax_rows(logits):
    """float32 softmax over dim 1 (reference: model.py:226)."""
    z = logits - logits.max(axis=1, keepdims=True)
    e = np.exp(z, dtype=np.float32)
    return (e / e.sum(axis=1, keepdims=True, dtype=np.float32)).astype(np.float32)


def predict_proba(x, weights, biases):
    """(probs f32 [nq,C] descending, classes i64 [nq,C]); reference: model.py:214-229.

    ``prob.topk(C)`` is a full descending sort; ties are broken arbitrarily by torch, here
    by lower class id first (stable sort on the negated probabilities).
    """
    p = softmax_rows(mlp_logits(x, weights, biases))
    order = np.argsort(-p, axis=1, kind="stable")
    return np.take_along_axis(p, order, axis=1), order.astype(np.int64)


def predict(x, weights, biases):
    """argmax of the logits = bucket of every object; reference: model.py:201-212."""
    return mlp_logits(x, weights, biases).argmax(axis=1).astype(np.int64)


# --------------------------------------------------------------------------------------
# distance (reference: search/li/utils.py:10-11 pairwise_cosine -> sklearn cosine_similarity)
# --------------------------------------------------------------------------------------
def normalize_rows64(a):
    """sklearn ``normalize(X, copy=True)`` on the float64 upcast: rows / max(norm, ->1 if 0)."""
    a = np.asarray(a).astype(np.float64)
    nrm = np.sqrt(np.einsum("ij,ij->i", a, a))
    nrm[nrm == 0.0] = 1.0
    return a / nrm[:, None]


def cosine_distance(x, y):
    """``1 - cosine_similarity(x, y)`` in float64; reference: utils.py:10-11, :19."""
    return 1.0 - normalize_rows64(x) @ normalize_rows64(y).T


# --------------------------------------------------------------------------------------
# faithful restatement of the rank / category loops (what the CPU baseline times)
# reference: LearnedIndex.py:22-101 (search), :103-195 (search_single), utils.py:14-43
# --------------------------------------------------------------------------------------
class BucketTable:
    """Object rows grouped by category (what ``groupby('category')`` yields each rank,
    LearnedIndex.py:143).  ``ids`` are the DataFrame labels = row + id_base (search.py:71-72),
    ascending inside every bucket like ``g.index``."""

    def __init__(self, pred_categories, id_base=1):
        pc = np.asarray(pred_categories).astype(np.int64)
        self.order = np.argsort(pc, kind="stable")
        self.cats, starts = np.unique(pc[self.order], return_index=True)
        self.bounds = np.append(starts, pc.shape[0])
        self.id_base = id_base

    @classmethod
    def single(cls, cat, rows, id_base=1):
        """A table holding ONE category whose objects are the given (ascending) dataset rows -- what a bounded CPU
        sample iterates when it times single categories of a large dataset."""
        t = cls.__new__(cls)
        t.order = np.asarray(rows, dtype=np.int64)
        t.cats = np.array([int(cat)])
        t.bounds = np.array([0, t.order.shape[0]])
        t.id_base = id_base
        return t

    def __iter__(self):
        for i, cat in enumerate(self.cats):
            yield int(cat), self.order[self.bounds[i]:self.bounds[i + 1]]


def _threshold_compact(result, threshold_rows, k):
    """utils.py:22-43: keep only the columns that hold at least one below-threshold
    distance (strict <), everything else 10_000.  Returns (None, None) when no column
    qualifies."""
    hit_r, hit_c = np.where(result < threshold_rows[:, None])          # utils.py:22-23
    cols = np.unique(hit_c)                                            # utils.py:26
    if cols.shape[0] == 0:
        return None, None                                              # utils.py:28-32
    width = max(cols.shape[0], k)                                      # utils.py:33
    out = np.full((result.shape[0], width), EMPTY_DIST, dtype=np.float64)
    out[hit_r, np.searchsorted(cols, hit_c)] = result[hit_r, hit_c]    # utils.py:35-42
    return out, cols


def search_single_loops(table, data_search, queries_search, category_per_query, k=10,
                        threshold_dist=None, shipped_padding=True, flagged=None):
    """One rank of the reference search; reference: LearnedIndex.py:103-195.

    ``shipped_padding=True`` reproduces the '<k objects' branch exactly as shipped
    (LearnedIndex.py:174-190, defect D1 of SURVEY.md: row 0 broadcast to every query of
    the category).  ``flagged`` (bool [nq]) receives the rows that went through it.
    """
    nq = queries_search.shape[0]
    nns = np.zeros((nq, k), dtype=np.uint32)                           # :137
    dists = np.full((nq, k), EMPTY_DIST, dtype=np.float64)             # :138
    category_per_query = np.asarray(category_per_query)
    for cat, rows in table:                                            # :143
        cat_idxs = np.where(category_per_query == cat)[0]              # :144
        if rows.shape[0] == 0 or cat_idxs.shape[0] == 0:               # :147
            continue
        obj_ids = rows + table.id_base                                 # g.index
        d = cosine_distance(queries_search[cat_idxs], data_search[rows])   # :166 / utils.py:19
        if threshold_dist is not None:                                 # :149-163
            d, cols = _threshold_compact(d, threshold_dist[cat_idxs], k)
            if d is None:
                continue
            obj_ids = obj_ids[cols]
        kk = k if k < d.shape[1] else d.shape[1]
        rel = np.argsort(d, axis=1, kind="quicksort")[:, :kk]          # :170-172
        if obj_ids.shape[0] < k:                                       # :174
            if flagged is not None:
                flagged[cat_idxs] = True
            if not shipped_padding:
                # intended semantics: write what exists, leave the rest (0, 10000.0)
                m = obj_ids.shape[0]
                rel_m = rel[:, :m]
                val = np.take_along_axis(d, rel_m, axis=1)
                real = (rel_m < m) & (val < EMPTY_DIST)
                got = np.where(real, obj_ids[np.minimum(rel_m, m - 1)], 0)
                nns[cat_idxs[:, None], np.arange(m)[None, :]] = got
                dists[cat_idxs[:, None], np.arange(m)[None, :]] = np.where(real, val, EMPTY_DIST)
                continue
            pad = (k - obj_ids.shape[0]) // 2 + 1                      # :176
            obj_ids = np.pad(obj_ids, pad, "edge")[:k]                 # :177-179
            rel = np.pad(rel[0], pad, "edge")[:k].reshape(1, -1)       # :180-182
            d = np.pad(d[0], pad, "edge")[:k].reshape(1, -1)           # :183-185
            _, first = np.unique(d, return_index=True)                 # :187
            d[0][np.setdiff1d(np.arange(k), first)] = EMPTY_DIST       # :188-190
        nns[cat_idxs] = obj_ids[rel]                                   # :192
        dists[cat_idxs] = np.take_along_axis(d, rel, axis=1)           # :193
    return dists, nns


def search_loops(weights, biases, queries_navigation, data_search, queries_search,
                 pred_categories, n_buckets=1, k=10, use_threshold=False, id_base=1,
                 shipped_padding=True, return_flags=False):
    """Restatement of ``LearnedIndex.search`` (LearnedIndex.py:22-101) with the same loop
    structure, cost profile and arithmetic as the reference: one float64 distance matrix
    per (rank, category), full-row argsort, stable rank merge.  This is the function the
    CPU baseline times (``bench.py`` cpu_baseline / ``--impl reference``)."""
    _, classes = predict_proba(queries_navigation, weights, biases)    # :61-63
    table = BucketTable(pred_categories, id_base)                      # :67 + :143
    data_search = np.asarray(data_search)
    queries_search = np.asarray(queries_search)
    flagged = np.zeros(queries_search.shape[0], dtype=bool)
    d_fin = a_fin = None
    for rank in range(n_buckets):                                      # :70
        thr = d_fin.max(axis=1) if (rank != 0 and use_threshold) else None   # :71-74
        d, a = search_single_loops(table, data_search, queries_search, classes[:, rank], k,
                                   thr, shipped_padding, flagged)      # :75-81
        if a_fin is None:
            a_fin, d_fin = a, d
        else:
            a_cat = np.hstack((a_fin, a))                              # :88
            d_cat = np.hstack((d_fin, d))                              # :89
            sel = np.argsort(d_cat, axis=1, kind="stable")[:, :k]      # :91
            d_fin = np.take_along_axis(d_cat, sel, axis=1)             # :94-97
            a_fin = np.take_along_axis(a_cat, sel, axis=1)
    if return_flags:
        return d_fin, a_fin, flagged
    return d_fin, a_fin


# --------------------------------------------------------------------------------------
# vectorised restatement of the intended semantics (SURVEY.md 8(a) semantics box)
# --------------------------------------------------------------------------------------
def search_intended(classes, data_search, queries_search, pred_categories, n_buckets, k=10,
                    id_base=1, thresholds=None):
    """result(q) = k smallest { 1 - cos64(q, x) : pred_categories[x] in classes[q, :n_buckets] }.

    Order: ascending distance; ties -> lower bucket rank first (strict '<' at utils.py:23 plus
    the stable merge at LearnedIndex.py:91), then ascending id.  A rank pointing at an empty
    category contributes nothing; unfilled slots are (id 0, dist 10000.0)
    (LearnedIndex.py:137-138).  ``thresholds`` (f64 [nq] or None): keep only distances
    strictly below it (single-rank use, LearnedIndex.py:149-163).

    Equal to the reference's own ``search(..., use_threshold=False)`` wherever every probed
    bucket holds >= k objects (checked by tests/test_oracle_golden.py).
    """
    classes = np.asarray(classes)[:, :n_buckets]
    nq = queries_search.shape[0]
    table = BucketTable(pred_categories, id_base)
    cand_d = [[] for _ in range(nq)]
    cand_key = [[] for _ in range(nq)]
    qn = normalize_rows64(queries_search)
    for cat, rows in table:
        qi, rank = np.where(classes == cat)
        if qi.shape[0] == 0:
            continue
        d = 1.0 - qn[qi] @ normalize_rows64(np.asarray(data_search)[rows]).T
        ids = rows + id_base
        kk = min(k, rows.shape[0])
        part = np.argpartition(d, kk - 1, axis=1)[:, :kk] if kk < rows.shape[0] else \
            np.tile(np.arange(rows.shape[0]), (qi.shape[0], 1))
        # keep everything tied with the kk-th value as well, so ties resolve by (rank, id)
        kth = np.take_along_axis(d, part, axis=1).max(axis=1)
        for j, q in enumerate(qi):
            sel = np.where(d[j] <= kth[j])[0]
            cand_d[q].append(d[j, sel])
            cand_key[q].append(np.stack([np.full(sel.shape[0], rank[j]), ids[sel]], axis=1))
    dists = np.full((nq, k), EMPTY_DIST, dtype=np.float64)
    nns = np.zeros((nq, k), dtype=np.uint32)
    for q in range(nq):
        if not cand_d[q]:
            continue
        d = np.concatenate(cand_d[q])
        key = np.concatenate(cand_key[q])
        if thresholds is not None:
            keep = d < thresholds[q]
            d, key = d[keep], key[keep]
        order = np.lexsort((key[:, 1], key[:, 0], d))[:k]
        dists[q, :order.shape[0]] = d[order]
        nns[q, :order.shape[0]] = key[order, 1]
    return dists, nns


def search_fast(weights, biases, queries_navigation, data_search, queries_search,
                pred_categories, n_buckets=1, k=10, id_base=1):
    _, classes = predict_proba(queries_navigation, weights, biases)
    return search_intended(classes, data_search, queries_search, pred_categories,
                           n_buckets, k, id_base)


# --------------------------------------------------------------------------------------
# ground truth and recall (reference: search/li/Baseline.py:14-20, eval/eval.py:37-45)
# --------------------------------------------------------------------------------------
def brute_force(queries, data, k=10, id_base=1, block=65536):
    """Exhaustive float64 cosine k-NN; reference: Baseline.py:17-19 (blocked over objects)."""
    nq = queries.shape[0]
    best_d = np.full((nq, 0), 0.0)
    best_i = np.zeros((nq, 0), dtype=np.int64)
    qn = normalize_rows64(queries)
    for s in range(0, data.shape[0], block):
        d = 1.0 - qn @ normalize_rows64(data[s:s + block]).T
        ids = np.arange(s, s + d.shape[1], dtype=np.int64)[None, :].repeat(nq, 0) + id_base
        d = np.hstack((best_d, d))
        ids = np.hstack((best_i, ids))
        sel = np.argsort(d, axis=1, kind="stable")[:, :k]
        best_d = np.take_along_axis(d, sel, axis=1)
        best_i = np.take_along_axis(ids, sel, axis=1)
    return best_d, best_i.astype(np.uint32)


def get_recall(found, truth, k=10):
    """reference: eval/eval.py:37-45 (set intersection per row; duplicates count once)."""
    assert k <= found.shape[1] and len(found) == len(truth)
    hit = 0
    for i in range(len(found)):
        hit += len(set(found[i, :k].tolist()) & set(truth[i, :k].tolist()))
    return hit / (len(found) * k)
