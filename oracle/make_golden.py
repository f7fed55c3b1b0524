"""Generate ``tests/golden/*.npz`` by executing the UNMODIFIED reference hot path.

Run in the build container (``/root/reference`` present):  ``python -m oracle.make_golden``

The reference has no tests, fixtures or golden vectors of its own, so
these files are the pin for both the oracle (tests/test_oracle_golden.py) and the CUDA path
(tests/test_gpu_parity.py): inputs + the exact arrays returned by the reference's own
``NeuralNetwork.predict_proba`` (model.py:214-229), ``LearnedIndex.search``
(LearnedIndex.py:22-101, both ``use_threshold`` settings), ``LearnedIndex.search_single``
(LearnedIndex.py:103-195) and ``Baseline.search`` (Baseline.py:14-20), run through
``oracle/ref_shims.py`` on this image's numpy/torch/scikit-learn (versions recorded in the
file).  ``/root/reference`` does not exist on the GPU box; the fixtures travel instead.
"""
from __future__ import annotations

import os
import sys

import numpy as np
import pandas as pd
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from oracle import ref_shims  # noqa: E402
import lmi_b200  # noqa: E402
from lmi_b200.synth import SynthSpec, make_synthetic  # noqa: E402

GOLDEN = os.path.join(ROOT, "tests", "golden")


def versions():
    import sklearn
    return dict(numpy=np.__version__, torch=torch.__version__, sklearn=sklearn.__version__,
                pandas=pd.__version__)


def params_of(li):
    lin = [m for m in li.model.model.layers if isinstance(m, torch.nn.Linear)]
    return ([m.weight.detach().numpy().copy() for m in lin], [m.bias.detach().numpy().copy() for m in lin])


def run_reference_case(tag, n, nq, n_categories, model_type, epochs, bucket_list, spec_kw, k=10):
    ref = ref_shims.load_reference()
    spec = SynthSpec(n=n, nq=nq, pca_rows=n, **spec_kw)
    arrs = {k_: v.numpy() for k_, v in make_synthetic(spec, "cpu").items()}
    data_nav = pd.DataFrame(arrs["data_nav"]); data_nav.index += 1          # search.py:71-72
    data_search = pd.DataFrame(arrs["data_f16"]); data_search.index += 1    # search.py:83-84
    q_nav, q_f16 = arrs["queries_nav"], arrs["queries_f16"]

    torch.manual_seed(2023); np.random.seed(2023)
    li = ref.index.LearnedIndex()
    pred_categories, _ = li.build(data_nav, n_categories=n_categories, epochs=epochs, lr=0.009,
                                  model_type=model_type)
    weights, biases = params_of(li)
    probs, classes = li.model.predict_proba(ref.model.data_X_to_torch(q_nav))

    out = dict(
        tag=tag, model_type=model_type, k=k, n_categories=n_categories,
        versions=str(versions()), spec=str(spec.as_dict()),
        data_f16=arrs["data_f16"], queries_f16=q_f16, queries_nav=q_nav,
        pred_categories=pred_categories.astype(np.int64),
        probs=probs, classes=classes.astype(np.int64),
        n_layers=len(weights), bucket_list=np.array(bucket_list),
    )
    for i, (w, b) in enumerate(zip(weights, biases)):
        out[f"w{i}"] = w; out[f"b{i}"] = b

    # single-bucket path exactly as run() does it (search.py:129-140)
    nav_copy = data_nav.copy(); nav_copy["category"] = pred_categories
    d1, a1 = li.search_single(data_navigation=nav_copy, data_search=data_search,
                              queries_search=q_f16, pred_categories=classes[:, 0], k=k)
    out["single_dists"], out["single_ids"] = d1, a1

    for nb in bucket_list:
        d, a = li.search(data_nav.copy(), q_nav, data_search, q_f16, pred_categories,
                         n_buckets=nb, k=k, use_threshold=False)
        out[f"clean_dists_nb{nb}"], out[f"clean_ids_nb{nb}"] = d, a
        with ref_shims.D1Flagger(nq, k) as fl:
            d, a = li.search(data_nav.copy(), q_nav, data_search, q_f16, pred_categories,
                             n_buckets=nb, k=k, use_threshold=True)
        out[f"shipped_dists_nb{nb}"], out[f"shipped_ids_nb{nb}"] = d, a
        out[f"shipped_d1_nb{nb}"] = fl.flags.copy()

    # the threshold path of search_single exactly as search() drives it for rank 1 (LearnedIndex.py:71-81,
    # :149-163 -> utils.py:14-43): categories = classes[:, 1], threshold = the k-th distance after rank 0.
    # Slots whose distance is 10000.0 carry arbitrary ids (utils.py:33-42 pads the compact matrix), rows that
    # went through the '<k' branch (D1) are flagged; consumers compare ids on slots with dist < 10000 of
    # unflagged rows.
    thr = d1.max(axis=1)
    with ref_shims.D1Flagger(nq, k) as fl:
        dt, at = li.search_single(data_navigation=nav_copy, data_search=data_search, queries_search=q_f16,
                                  pred_categories=classes[:, 1], k=k, threshold_dist=thr)
    out["single_thr"], out["single_thr_dists"], out["single_thr_ids"] = thr, dt, at
    out["single_thr_d1"] = fl.flags.copy()
    # a tighter threshold (median of the rank-0 distances) on rank 0's own categories: most slots are pruned
    thr2 = np.median(d1, axis=1)
    with ref_shims.D1Flagger(nq, k) as fl2:
        dt2, at2 = li.search_single(data_navigation=nav_copy, data_search=data_search, queries_search=q_f16,
                                    pred_categories=classes[:, 0], k=k, threshold_dist=thr2)
    out["single_thr2"], out["single_thr2_dists"], out["single_thr2_ids"] = thr2, dt2, at2
    out["single_thr2_d1"] = fl2.flags.copy()

    # brute-force ground truth from the reference's own Baseline (Baseline.py:14-20)
    sub = min(64, nq)
    bd, bi, _ = ref.baseline.Baseline().search(queries=q_f16[:sub], data=arrs["data_f16"], k=k)
    out["baseline_dists"], out["baseline_ids"] = bd, bi.astype(np.uint32)

    path = os.path.join(GOLDEN, f"{tag}.npz")
    np.savez_compressed(path, **out)
    sizes = np.bincount(pred_categories, minlength=n_categories)
    print(f"{tag}: wrote {path} ({os.path.getsize(path)/1e6:.2f} MB); bucket sizes min/med/max "
          f"{sizes.min()}/{int(np.median(sizes))}/{sizes.max()}, empty={int((sizes == 0).sum())}, "
          f"<k={int(((sizes > 0) & (sizes < k)).sum())}")


def main():
    os.makedirs(GOLDEN, exist_ok=True)
    blobs = dict(n_blobs=80, sigma=0.7)
    # what search.py actually trains (model_type never forwarded -> 'MLP')
    run_reference_case("ref_mlp_n2400_c24", 2400, 256, 24, "MLP", 40, [2, 5], blobs)
    # the README / BASELINE architecture (96->256->128->C)
    run_reference_case("ref_mlp5_n1600_c16", 1600, 192, 16, "MLP-5", 40, [3], blobs)
    # many categories, few objects: empty buckets and buckets with < k objects (defect D1 territory)
    run_reference_case("ref_mlp_tiny_buckets", 900, 128, 40, "MLP", 60, [4], dict(n_blobs=48, sigma=0.5))


if __name__ == "__main__":
    main()
