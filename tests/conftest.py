import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN_DIR = os.path.join(ROOT, "tests", "golden")
GOLDEN_TAGS = ["ref_mlp_n2400_c24", "ref_mlp5_n1600_c16", "ref_mlp_tiny_buckets"]


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def load_golden(tag):
    z = np.load(os.path.join(GOLDEN_DIR, tag + ".npz"), allow_pickle=False)
    g = {k: z[k] for k in z.files}
    n_layers = int(g["n_layers"])
    g["weights"] = [g[f"w{i}"] for i in range(n_layers)]
    g["biases"] = [g[f"b{i}"] for i in range(n_layers)]
    g["k"] = int(g["k"])
    g["n_categories"] = int(g["n_categories"])
    g["bucket_list"] = [int(b) for b in g["bucket_list"]]
    return g


@pytest.fixture(params=GOLDEN_TAGS)
def golden(request):
    return load_golden(request.param)


def rows_with_small_buckets(g, nb):
    """Query rows that probe a bucket holding 0 < n < k objects within their first nb ranks
    (where the reference's '<k' padding branch, LearnedIndex.py:174-190, fires on the
    no-threshold path)."""
    sizes = np.bincount(g["pred_categories"], minlength=g["n_categories"])
    small = (sizes > 0) & (sizes < g["k"])
    return small[g["classes"][:, :nb]].any(axis=1)


def assert_ids_equal_up_to_ties(ids, dists, ref_ids, ref_dists, tie_eps=1e-12, what=""):
    """Bit-exact ids, except inside groups of reference distances that differ by <= tie_eps
    (documented equal-distance ties): there the id *sets* must agree."""
    assert ids.shape == ref_ids.shape, what
    bad = np.where((ids != ref_ids).any(axis=1))[0]
    for r in bad:
        d = ref_dists[r]
        cols = np.where(ids[r] != ref_ids[r])[0]
        for c in cols:
            grp = np.where(np.abs(d - d[c]) <= tie_eps)[0]
            assert set(ids[r, grp].tolist()) == set(ref_ids[r, grp].tolist()), \
                f"{what}: row {r} col {c}: ids {ids[r]} vs reference {ref_ids[r]} (dists {d})"


# name(s) the default tcgen05 kernel may report: LMI_FORCE_PAIR_KERNEL=1 reroutes every default-kernel caller to the
# CTA-pair kernel (lmi_b200.index), which is how the whole GPU suite is run against it
import os as _os
TS_KERNELS = ("rescore_ts2_kernel",) if _os.environ.get("LMI_FORCE_PAIR_KERNEL", "0") == "1" else ("rescore_ts_kernel",)
