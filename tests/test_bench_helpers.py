"""bench.py's in-run checkers are themselves checked against the oracle on the golden fixtures (CPU)."""
import os
import sys

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
for p in (ROOT, HERE):
    if p not in sys.path:
        sys.path.insert(0, p)

from conftest import load_golden, assert_ids_equal_up_to_ties  # noqa: E402


@pytest.mark.parametrize("nb", [1, 3])
def test_bucket_restricted_truth_equals_oracle(nb):
    import torch
    import bench
    from oracle import lmi_oracle as O
    g = load_golden("ref_mlp_n2400_c24")
    k = g["k"]
    od, oa = O.search_intended(g["classes"], g["data_f16"], g["queries_f16"], g["pred_categories"], nb, k)
    sel = np.arange(0, g["queries_f16"].shape[0], 7)
    d, i = bench.bucket_restricted_truth(torch.from_numpy(g["data_f16"]), torch.from_numpy(g["pred_categories"].astype(np.int64)), 1,
                                         torch.from_numpy(g["queries_f16"][sel]),
                                         torch.from_numpy(g["classes"][sel, :nb].astype(np.int64)), k)
    assert_ids_equal_up_to_ties(i.numpy().astype(np.uint32), d.numpy(), oa[sel], od[sel], what="bench checker")
    np.testing.assert_allclose(d.numpy(), od[sel], rtol=0, atol=1e-12)


def test_ground_truth_returns_sorted_dists_with_id_base():
    import torch
    import bench
    g = load_golden("ref_mlp_n2400_c24")
    x, q = torch.from_numpy(g["data_f16"]), torch.from_numpy(g["queries_f16"][:20])
    ids = bench.ground_truth(x, q, 10)
    d2, ids2 = bench.ground_truth(x, q, 10, id_base=101, return_dists=True)
    assert torch.equal(ids + 100, ids2) and bool((d2[:, 1:] >= d2[:, :-1]).all())
    assert bench.recall_at_k(ids, torch.from_numpy(g["baseline_ids"][:20].astype(np.int64)), 10) > 0.999
