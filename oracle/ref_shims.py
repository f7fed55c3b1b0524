"""Import the UNMODIFIED reference (``/root/reference/search/li``) in the build container.

TEST INFRASTRUCTURE ONLY.  Used by ``oracle/make_golden.py`` (fixture generation) and by
``tests/test_oracle_vs_reference.py`` (skipped when ``/root/reference`` is absent, e.g. on
the GPU box).  No reference file is edited or copied; three import-time shims make the hot
path run on this image:

1. ``faiss`` is not installed  -> stub module with a Lloyd ``Kmeans`` (only labels matter;
   used by ``LearnedIndex.cluster``, LearnedIndex.py:273-282, i.e. index build only).
2. ``h5py`` is not installed   -> empty stub (used only by I/O helpers, utils.py:85-97).
3. numpy >= 2: ``np.ogrid[...]`` returns a tuple and LearnedIndex.py:95 assigns into it
   -> the module-level name ``np`` inside ``li.LearnedIndex`` is replaced by a proxy whose
   ``ogrid[...]`` returns a list.
"""
from __future__ import annotations

import os
import sys
import types

import numpy as np

REFERENCE_ROOT = os.environ.get("LMI_REFERENCE_ROOT", "/root/reference")


def reference_available() -> bool:
    return os.path.isdir(os.path.join(REFERENCE_ROOT, "search", "li"))


class _LloydIndex:
    def __init__(self, centroids):
        self.c = centroids

    def search(self, x, k):
        assert k == 1
        d = (x * x).sum(1)[:, None] - 2.0 * x @ self.c.T + (self.c * self.c).sum(1)[None, :]
        lab = d.argmin(1)
        return d[np.arange(x.shape[0]), lab][:, None], lab[:, None]


class _LloydKmeans:
    """Stand-in for ``faiss.Kmeans(d, k, verbose, seed)``: ``train(X)`` then ``index.search(X, 1)``."""

    def __init__(self, d, k, verbose=False, seed=0, niter=15):
        self.d, self.k, self.seed, self.niter = d, k, seed, niter
        self.index = None

    def train(self, x):
        rng = np.random.default_rng(self.seed)
        c = x[rng.choice(x.shape[0], self.k, replace=False)].astype(np.float32).copy()
        for _ in range(self.niter):
            lab = _LloydIndex(c).search(x, 1)[1][:, 0]
            for j in range(self.k):
                m = lab == j
                if m.any():
                    c[j] = x[m].mean(0)
        self.index = _LloydIndex(c)


class _OgridList:
    def __getitem__(self, item):
        return list(np.ogrid[item])


class _NumpyProxy:
    ogrid = _OgridList()

    def __getattr__(self, name):
        return getattr(np, name)


_loaded = None


def load_reference():
    """Returns the reference modules (LearnedIndex, model, utils, Baseline), MLP pinned to CPU."""
    global _loaded
    if _loaded is not None:
        return _loaded
    if not reference_available():
        raise RuntimeError(f"reference not found under {REFERENCE_ROOT}")
    if "faiss" not in sys.modules:
        faiss = types.ModuleType("faiss")
        faiss.Kmeans = _LloydKmeans
        sys.modules["faiss"] = faiss
    if "h5py" not in sys.modules:
        try:
            import h5py  # noqa: F401
        except Exception:
            sys.modules["h5py"] = types.ModuleType("h5py")
    sys.path.insert(0, os.path.join(REFERENCE_ROOT, "search"))
    import li.LearnedIndex as ref_index
    import li.model as ref_model
    import li.utils as ref_utils
    import li.Baseline as ref_baseline
    import torch
    ref_index.np = _NumpyProxy()
    ref_model.get_device = lambda: torch.device("cpu")   # model.py:108-109 would pick cuda:0
    ref_index.tqdm = lambda it, *a, **k: it               # silence the progress bar
    _loaded = types.SimpleNamespace(index=ref_index, model=ref_model, utils=ref_utils,
                                    baseline=ref_baseline)
    return _loaded


class D1Flagger:
    """Context manager recording which query rows pass through the reference's '<k objects'
    padding branch (LearnedIndex.py:174-190, defect D1 in SURVEY.md 0.4) on the threshold
    path.  (On the no-threshold path the branch fires iff a probed bucket holds fewer than
    k objects, which callers derive from ``pred_categories`` directly.)"""

    def __init__(self, nq, k=10):
        self.flags = np.zeros(nq, dtype=bool)
        self.k = k

    def __enter__(self):
        ref = load_reference()
        self._ref = ref
        self._orig_thr = ref.index.pairwise_cosine_threshold

        def thr(x, y, threshold, cat_idxs, k=10):
            out = self._orig_thr(x, y, threshold, cat_idxs, k)
            if len(out) == 3 and len(out[1]) < self.k:
                self.flags[cat_idxs] = True
            return out

        ref.index.pairwise_cosine_threshold = thr
        return self

    def __exit__(self, *exc):
        self._ref.index.pairwise_cosine_threshold = self._orig_thr
        return False


def make_reference_index(weights, biases, model_type):
    """An unmodified ``li.LearnedIndex`` whose model carries the given weights
    (what ``build`` would leave in ``self.model``, LearnedIndex.py:231-239)."""
    import torch
    ref = load_reference()
    in_dim = weights[0].shape[1]
    out_dim = weights[-1].shape[0]
    nn = ref.model.NeuralNetwork(input_dim=in_dim, output_dim=out_dim, lr=0.01, model_type=model_type)
    linear = [m for m in nn.model.layers if isinstance(m, torch.nn.Linear)]
    assert len(linear) == len(weights)
    with torch.no_grad():
        for m, w, b in zip(linear, weights, biases):
            m.weight.copy_(torch.from_numpy(np.asarray(w, dtype=np.float32)))
            m.bias.copy_(torch.from_numpy(np.asarray(b, dtype=np.float32)))
    li = ref.index.LearnedIndex()
    li.model = nn
    return li
