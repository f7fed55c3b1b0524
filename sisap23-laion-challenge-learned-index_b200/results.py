"""Result files and recall of the query path.

``store_results`` keeps the schema the reference writes (search/li/utils.py:85-97) and the SISAP
evaluator reads (eval/eval.py:64-76): datasets ``knns`` (uint32 [nq, k], 1-based ids) and ``dists``
(float64 [nq, k]) plus the attributes ``algo, data, buildtime, querytime, size, params``.  HDF5 is
written when ``h5py`` is importable (it is not on the build image); otherwise -- or when the
destination ends in ``.npz`` -- the same fields go into an ``.npz`` (attributes as 0-d arrays, the
layout ``eval.py``'s ``return_h5_str`` fallback reads: ``res["size"][()]``).  ``load_results`` reads
either.  ``get_recall`` restates eval/eval.py:37-45 (set intersection per row, duplicates collapse).
"""
from __future__ import annotations

import os
from pathlib import Path

import numpy as np

ATTRS = ("algo", "data", "buildtime", "querytime", "size", "params")


def _have_h5py():
    try:
        import h5py
        if not hasattr(h5py, "File"):          # an empty stand-in module (e.g. oracle/ref_shims.py's import shim)
            return False
        return True
    except Exception:
        return False


def result_path(dst):
    """The file ``store_results(dst, ...)`` will write: ``dst`` itself for HDF5, else ``dst`` with
    the suffix replaced by ``.npz``."""
    dst = str(dst)
    if dst.endswith(".npz") or not _have_h5py():
        return str(Path(dst).with_suffix(".npz"))
    return dst


def store_results(dst, algo, kind, dists, anns, buildtime, querytime, params, size):
    """Same signature and fields as the reference's ``store_results`` (utils.py:85-97)."""
    dists, anns = np.asarray(dists), np.asarray(anns)
    if dists.shape != anns.shape or dists.ndim != 2:
        raise ValueError(f"dists {dists.shape} and anns {anns.shape} must be equal 2-d shapes")
    path = result_path(dst)
    os.makedirs(Path(path).parent, exist_ok=True)
    attrs = dict(algo=algo, data=kind, buildtime=buildtime, querytime=querytime, size=size, params=params)
    if path.endswith(".npz"):
        with open(path, "wb") as fh:           # a file object: numpy must not append a second suffix
            np.savez(fh, knns=anns, dists=dists, **{k: np.asarray(v) for k, v in attrs.items()})
        return path
    import h5py
    with h5py.File(path, "w") as f:
        for k, v in attrs.items():
            f.attrs[k] = v
        f.create_dataset("knns", anns.shape, dtype=anns.dtype)[:] = anns
        f.create_dataset("dists", dists.shape, dtype=dists.dtype)[:] = dists
    return path


def load_results(path):
    """-> dict with ``knns``, ``dists`` and the six attributes (python scalars / str)."""
    path = str(path)
    out = {}
    if path.endswith(".npz"):
        with np.load(path, allow_pickle=False) as z:
            out["knns"], out["dists"] = z["knns"], z["dists"]
            for k in ATTRS:
                v = z[k][()]
                out[k] = v.item() if isinstance(v, np.generic) else v
        return out
    import h5py
    with h5py.File(path, "r") as f:
        out["knns"], out["dists"] = np.array(f["knns"]), np.array(f["dists"])
        for k in ATTRS:
            v = f.attrs[k]
            out[k] = v.decode() if isinstance(v, bytes) else (v.item() if isinstance(v, np.generic) else v)
    return out


def get_recall(I, gt, k):
    """eval/eval.py:37-45: mean over queries of |set(I[i,:k]) & set(gt[i,:k])| / k."""
    I, gt = np.asarray(I), np.asarray(gt)
    assert k <= I.shape[1]
    assert len(I) == len(gt)
    n = len(I)
    hits = 0
    for i in range(n):
        hits += len(set(I[i, :k].tolist()) & set(gt[i, :k].tolist()))
    return hits / (n * k)


def get_recall_fast(I, gt, k):
    """Vectorised ``get_recall`` for 10k x 10 .. 10k x 100 (sort + searchsorted per row would be
    overkill: a broadcast compare is 1e7 booleans at k = 100).  Identical value: duplicates inside a
    row of ``I`` or ``gt`` are counted once, as the set semantics require."""
    I, gt = np.asarray(I)[:, :k], np.asarray(gt)[:, :k]
    assert len(I) == len(gt)
    Is = np.sort(I, axis=1)
    first = np.ones(Is.shape, dtype=bool)
    first[:, 1:] = Is[:, 1:] != Is[:, :-1]                     # one representative per distinct id
    in_gt = (Is[:, :, None] == gt[:, None, :]).any(axis=2)
    return float((in_gt & first).sum()) / (len(I) * k)
