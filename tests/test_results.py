"""N3: result files keep the reference schema (search/li/utils.py:85-97) and
``get_recall`` equals eval/eval.py:37-45."""
import os
import sys

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
for p in (ROOT, HERE):
    if p not in sys.path:
        sys.path.insert(0, p)

from lmi_b200 import results as R  # noqa: E402


def _eval_get_recall(I, gt, k):
    # the evaluator's loop, restated here independently (eval/eval.py:37-45)
    assert k <= I.shape[1] and len(I) == len(gt)
    rec = 0
    for i in range(len(I)):
        rec += len(set(I[i, :k]) & set(gt[i, :k]))
    return rec / (len(I) * k)


def test_store_and_load_roundtrip(tmp_path):
    rng = np.random.default_rng(0)
    anns = rng.integers(1, 1000, size=(37, 10)).astype(np.uint32)
    dists = np.sort(rng.random((37, 10)), axis=1)
    dst = tmp_path / "result" / "clip768v2" / "300K" / "run.h5"
    path = R.store_results(str(dst), "LMI", "clip768v2", dists, anns, 12.5, 0.25, "bp=4", "300K")
    assert os.path.exists(path) and path == R.result_path(str(dst))
    z = R.load_results(path)
    assert z["knns"].dtype == np.uint32 and z["dists"].dtype == np.float64
    assert np.array_equal(z["knns"], anns) and np.array_equal(z["dists"], dists)
    assert (z["algo"], z["data"], z["size"], z["params"]) == ("LMI", "clip768v2", "300K", "bp=4")
    assert z["buildtime"] == 12.5 and z["querytime"] == 0.25


def test_npz_fields_read_like_the_evaluator_fallback(tmp_path):
    # eval/eval.py:47-53,64-66 falls back to res[param][()] when attrs are absent
    anns = np.arange(20, dtype=np.uint32).reshape(2, 10) + 1
    path = R.store_results(str(tmp_path / "r.npz"), "LMI", "clip768v2", anns.astype(np.float64), anns, 1.0, 2.0, "p", "10M")
    with np.load(path) as z:
        assert str(z["size"][()]) == "10M" and float(z["querytime"][()]) == 2.0
        assert set(z.files) == {"knns", "dists", *R.ATTRS}


def test_shape_mismatch_is_rejected(tmp_path):
    with pytest.raises(ValueError):
        R.store_results(str(tmp_path / "x.npz"), "a", "b", np.zeros((3, 10)), np.zeros((3, 9), np.uint32), 0, 0, "", "")


@pytest.mark.parametrize("k", [1, 10, 30])
def test_get_recall_matches_evaluator(k):
    rng = np.random.default_rng(k)
    gt = np.stack([rng.permutation(200)[:30] for _ in range(64)]).astype(np.uint32) + 1
    I = gt.copy()
    I[rng.random(I.shape) < 0.3] = 0                      # unfilled slots (id 0), duplicates inside a row
    I[::7, :5] = I[::7, 5:10]                             # duplicated true neighbours count once
    want = _eval_get_recall(I, gt, k)
    assert R.get_recall(I, gt, k) == want
    assert abs(R.get_recall_fast(I, gt, k) - want) < 1e-15


def test_get_recall_asserts_like_the_evaluator():
    with pytest.raises(AssertionError):
        R.get_recall(np.zeros((2, 5)), np.zeros((2, 10)), 10)
    with pytest.raises(AssertionError):
        R.get_recall(np.zeros((2, 10)), np.zeros((3, 10)), 10)


@pytest.mark.skipif(not os.path.exists("/root/reference/eval/eval.py"), reason="reference checkout not present on this machine")
def test_get_recall_equals_reference_function():
    """Where the reference checkout exists (the build container), call its own get_recall."""
    import importlib.util
    import types
    stub = {m: sys.modules.get(m) for m in ("h5py",)}
    if stub["h5py"] is None:
        sys.modules["h5py"] = types.ModuleType("h5py")           # eval.py imports it at module level only for I/O
    try:
        spec = importlib.util.spec_from_file_location("_ref_eval", "/root/reference/eval/eval.py")
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
    finally:
        if stub["h5py"] is None:
            del sys.modules["h5py"]
    rng = np.random.default_rng(5)
    gt = np.stack([rng.permutation(500)[:10] for _ in range(100)]).astype(np.uint32) + 1
    I = np.where(rng.random(gt.shape) < 0.2, rng.integers(1, 500, gt.shape), gt).astype(np.uint32)
    assert R.get_recall(I, gt, 10) == mod.get_recall(I, gt, 10)
    assert abs(R.get_recall_fast(I, gt, 10) - mod.get_recall(I, gt, 10)) < 1e-15
