"""Calibrates bench.py's CPU arm against the UNMODIFIED reference.

Config 0 of BASELINE.json at its real shape -- 300K objects, C = 122, `-bp 6` -> 7 leaf buckets (search.py:37-38),
10k queries, k = 10 -- on synthetic LAION-shaped data, in the build container (the reference does not travel to the GPU
box).  Times, on the same inputs and the same weights:

  reference : ``li.search(data, queries, data_search, queries_search, pred_categories, n_buckets=7, k=10,
              use_threshold=True)`` exactly as search/search.py:119-128 calls it (imported in place through
              oracle/ref_shims.py, no reference file edited), wall clock around the call like search.py:116,141
  port      : ``oracle.search_loops(..., use_threshold=True)`` -- what bench.py's cpu_baseline / --impl reference time
  estimator : bench.py's bounded sample (a few categories, real thresholds, scaled by pair fraction) on the same inputs

    python scripts/calibrate_cpu_baseline.py [--threads N] [--rows 300000] [--out profiles/r2_cpu_calibration.json]

Run once with all cores and once with --threads 1 (comparable with README.md:32-38: 22.08 s on one EPYC 7532 core).
"""
import argparse
import json
import os
import sys
import time

ap = argparse.ArgumentParser()
ap.add_argument("--threads", type=int, default=0, help="0 = all cores")
ap.add_argument("--rows", type=int, default=300_000)
ap.add_argument("--nq", type=int, default=10_000)
ap.add_argument("--buckets", type=int, default=7)
ap.add_argument("--out", default=None)
a = ap.parse_args()
if a.threads:
    for v in ("OMP_NUM_THREADS", "OPENBLAS_NUM_THREADS", "MKL_NUM_THREADS"):
        os.environ[v] = str(a.threads)

import numpy as np
import pandas as pd
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
if a.threads:
    torch.set_num_threads(a.threads)
from oracle import lmi_oracle as O, ref_shims          # noqa: E402
import lmi_b200                                         # noqa: E402,F401
from lmi_b200.synth import SynthSpec, make_synthetic    # noqa: E402
from lmi_b200.trainer import train_index                # noqa: E402
import bench                                            # noqa: E402

C, k, nb = 122, 10, a.buckets
spec = SynthSpec(n=a.rows, nq=a.nq, n_blobs=2000, sigma=0.55)
t = make_synthetic(spec, "cpu")
_, w, b = train_index(t["data_nav"], C, "MLP-5", sample=200_000, epochs=3)
arr = {kk: v.numpy() for kk, v in t.items()}
pc = O.predict(arr["data_nav"], w, b)

ref = ref_shims.load_reference()
li = ref_shims.make_reference_index(w, b, "MLP-5")
data = pd.DataFrame(arr["data_nav"]); data.index += 1                   # search.py:71-72
data_search = pd.DataFrame(arr["data_f16"]); data_search.index += 1    # search.py:83-84
out = dict(config=f"{a.rows} x 768 fp16, C={C}, MLP-5, n_buckets={nb}, nq={a.nq}, k={k}", threads=a.threads or os.cpu_count(),
           host_cpu_count=os.cpu_count())

s = time.time()
d_ref, a_ref = li.search(data_navigation=data, queries_navigation=arr["queries_nav"], data_search=data_search,
                         queries_search=arr["queries_f16"], pred_categories=pc, n_buckets=nb, k=k, use_threshold=True)
out["reference_seconds"] = time.time() - s
print("reference", out["reference_seconds"], flush=True)

# the port on the same memory layout as the reference's data_search (a DataFrame block is feature-major), so the
# per-call bucket gather (LearnedIndex.py:152/168) costs what it costs there; row-major timing beside it
data_fm = np.asfortranarray(arr["data_f16"])
s = time.time()
d_port, a_port = O.search_loops(w, b, arr["queries_nav"], data_fm, arr["queries_f16"], pc, n_buckets=nb, k=k,
                                use_threshold=True)
out["port_seconds"] = time.time() - s
s = time.time()
O.search_loops(w, b, arr["queries_nav"], arr["data_f16"], arr["queries_f16"], pc, n_buckets=nb, k=k, use_threshold=True)
out["port_seconds_row_major_data"] = time.time() - s
out["port_over_reference"] = out["port_seconds"] / out["reference_seconds"]
out["port_rows_identical_to_reference"] = float((a_port == a_ref).all(axis=1).mean())
print("port", out["port_seconds"], flush=True)


class Args:
    pass


args = Args()
args.buckets, args.k, args.categories, args.nq = nb, k, C, a.nq
inp = dict(queries_nav=t["queries_nav"], queries_f16=t["queries_f16"], data_f16=t["data_f16"], weights=w, biases=b,
           pred_categories=torch.from_numpy(pc))
inp["pred_categories"] = torch.from_numpy(pc)
est = bench.cpu_reference_sample(inp, args, 8.0)
out["estimator_seconds_full"] = est["est_seconds_full"]
out["estimator_over_port"] = est["est_seconds_full"] / out["port_seconds"]
out["estimator_over_reference"] = est["est_seconds_full"] / out["reference_seconds"]
out["estimator_sample"] = est["sample"]
out["qps"] = dict(reference=a.nq / out["reference_seconds"], port=a.nq / out["port_seconds"], estimator=est["value"])
print(json.dumps(out, indent=1))
if a.out:
    json.dump(out, open(a.out, "w"), indent=1)
