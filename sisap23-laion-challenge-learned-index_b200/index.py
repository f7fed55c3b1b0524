"""Host-side mirror of the reference's index interface for the query-time hot path.

``GpuLearnedIndex`` keeps the names, argument meaning, return types and error behaviour of
``li.LearnedIndex`` (reference: search/li/LearnedIndex.py:17-195) for ``search`` and
``search_single``, and ``GpuNavigator`` those of ``li.model.NeuralNetwork.predict_proba`` /
``predict`` (search/li/model.py:201-229).  All compute happens in liblmi_b200.so (sm_100a CUDA)
through the C-ABI of include/lmi_b200.h; there is no CPU path.

Semantics implemented: for each query the ``n_buckets`` most probable
categories of the MLP; candidates = every object whose ``pred_categories`` entry is one of them;
result = the k smallest ``1 - cos`` (float64 on the float16 inputs) with 1-based ids, ascending,
ties -> lower bucket rank then lower id; unfilled slots (id 0, dist 10000.0).  That is what the
reference's ``search(..., use_threshold=False)`` returns, and what ``use_threshold=True`` returns
on every row its '<k objects' padding defect (LearnedIndex.py:174-190) does not touch;
``use_threshold`` is accepted and ignored because the threshold is a pruning device, not part of
the result definition.
"""
from __future__ import annotations

import ctypes as C
import hashlib
import os
import struct
import weakref

import numpy as np

from . import _lib
from ._lib import LmiStats, check

RESCORE_SIMT, RESCORE_TC, RESCORE_TC_SS, RESCORE_TC_PAIR = 0, 1, 2, 3
SHARD_BUCKETS, SHARD_STRIPES = 0, 1


def _ptr(a):
    return C.c_void_p(a.ctypes.data) if a is not None else C.c_void_p(0)


def _view_rows(x):
    """DataFrame / ndarray -> (2-D ndarray VIEW, id_base) without copying the values.  The reference indexes objects by
    the DataFrame label, which run() makes 1-based (search.py:71-72, :83-84); a 'category' column is dropped like
    LearnedIndex.py:140-141.  A homogeneous frame hands out its block as it lies -- for a frame built from a 2-D array
    that is an F-ordered (feature-major) view, which ``lmi_store_append_strided`` uploads without a host transpose."""
    id_base = 1
    if hasattr(x, "to_numpy") and hasattr(x, "index"):
        idx = x.index
        n = len(idx)
        if hasattr(idx, "start") and hasattr(idx, "step"):                  # RangeIndex: O(1)
            if n > 1 and idx.step != 1:
                raise ValueError("data_search index must be consecutive integers")
            id_base = int(idx.start) if n else 1
        elif n:
            lab = np.asarray(idx)
            if not np.array_equal(lab, np.arange(lab[0], lab[0] + n)):
                raise ValueError("data_search index must be consecutive integers")
            id_base = int(lab[0])
        if "category" in x.columns:
            x = x.drop(columns="category")                                  # LearnedIndex.py:140-141 (a copy, like there)
        x = x.to_numpy()
    return np.asarray(x), id_base


def _as_rows(x):
    """Same, as a C-contiguous array (small inputs: query batches, GpuBaseline)."""
    rows, id_base = _view_rows(x)
    return np.ascontiguousarray(rows), id_base


def _alive(ref, obj):
    return ref is not None and ref() is obj


def _weak(obj):
    try:
        return weakref.ref(obj)
    except TypeError:
        return None


def extract_mlp(model):
    """(weights, biases) of the nn.Linear layers of a reference model: accepts li.LearnedIndex,
    li.model.NeuralNetwork, li.model.Model or a torch Sequential (model.py:15-83)."""
    import torch
    m = model
    for attr in ("model", "model", "layers"):
        if not isinstance(m, torch.nn.Sequential) and hasattr(m, attr):
            m = getattr(m, attr)
    mods = list(m) if isinstance(m, torch.nn.Sequential) else list(m.modules())
    lin = [x for x in mods if isinstance(x, torch.nn.Linear)]
    if not lin:
        raise ValueError("no nn.Linear layers found")
    for x in mods:
        if not isinstance(x, (torch.nn.Linear, torch.nn.ReLU, torch.nn.Sequential)):
            raise ValueError(f"unsupported layer {type(x).__name__}: only Linear/ReLU stacks (model.py:21-78)")
    w = [x.weight.detach().cpu().numpy().astype(np.float32) for x in lin]
    b = [x.bias.detach().cpu().numpy().astype(np.float32) for x in lin]
    return w, b


class GpuNavigator:
    """Stands where ``li.model.NeuralNetwork`` stands at query time (model.py:201-229)."""

    def __init__(self, owner: "GpuLearnedIndex"):
        self._o = owner

    def predict_proba(self, data_X, n_top=None):
        """-> (probs f32 [nq,n_top] descending, classes i64 [nq,n_top]); n_top defaults to all C
        categories like ``prob.topk(prob.shape[1])`` (model.py:226-229)."""
        o = self._o
        x = np.ascontiguousarray(np.asarray(data_X, dtype=np.float32))
        if x.ndim != 2 or x.shape[1] != o.d_nav:
            raise ValueError(f"expected [n, {o.d_nav}] navigation vectors")
        n_top = o.n_categories if n_top is None else int(n_top)
        classes = np.empty((x.shape[0], n_top), dtype=np.int64)
        probs = np.empty((x.shape[0], n_top), dtype=np.float32)
        check(o._lib.lmi_navigate(o._h, _ptr(x), x.shape[0], n_top, _ptr(classes), _ptr(probs)))
        return probs, classes

    def predict(self, data_X):
        """argmax category per row, int64 (model.py:201-212)."""
        o = self._o
        x = np.ascontiguousarray(np.asarray(data_X, dtype=np.float32))
        out = np.empty(x.shape[0], dtype=np.int64)
        check(o._lib.lmi_predict(o._h, _ptr(x), x.shape[0], _ptr(out)))
        return out


class GpuLearnedIndex:
    """Drop-in for ``li.LearnedIndex`` on the query path (LearnedIndex.py:17-195)."""

    def __init__(self, n_categories, d_nav=96, d=768, device=0, rescore=None,
                 chunk_objects=None, rank=0, world=1, shard_policy=SHARD_BUCKETS):
        self._lib = _lib.load()
        self._h = C.c_void_p()
        self.n_categories, self.d_nav, self.d, self.device = int(n_categories), int(d_nav), int(d), int(device)
        check(self._lib.lmi_create(C.byref(self._h), self.device, self.d_nav, self.d, self.n_categories))
        if rescore is None:
            # default: the CTA-pair tcgen05 kernel, sharded or not; the library falls back
            # to the smem-operand / CUDA-core kernels for shapes it does not cover (d > 768).  LMI_RESCORE=<0..3> overrides.
            env = os.environ.get("LMI_RESCORE", "")
            if env != "" and env not in ("0", "1", "2", "3"):
                raise ValueError(f"LMI_RESCORE must be 0..3, got {env!r}")
            rescore = int(env) if env != "" else RESCORE_TC_PAIR
        if rescore == RESCORE_TC and os.environ.get("LMI_FORCE_PAIR_KERNEL", "0") == "1":
            rescore = RESCORE_TC_PAIR            # test aid: run every single-CTA-kernel caller on the CTA-pair kernel
        self.set_option("rescore", rescore)
        if chunk_objects is not None:
            self.set_option("chunk_objects", chunk_objects)
        env_sets = os.environ.get("LMI_EPI_SETS", "")          # CTA-pair kernel: epilogue warp sets (0 = automatic, 1, 2)
        if env_sets != "":
            if env_sets not in ("0", "1", "2"):
                raise ValueError(f"LMI_EPI_SETS must be 0, 1 or 2, got {env_sets!r}")
            self.set_option("epi_sets", int(env_sets))
        for item in filter(None, os.environ.get("LMI_OPTIONS", "").split(",")):      # development: "name=value,..." -> lmi_set_option
            name, _, value = item.partition("=")
            self.set_option(name.strip(), int(value))
        if world > 1:
            check(self._lib.lmi_set_shard(self._h, rank, world, shard_policy))
        self.rank, self.world = rank, world
        self.shard_policy = int(shard_policy)
        self.model = None                 # like LearnedIndex.model (LearnedIndex.py:19-20)
        self._store_key = None            # (content key, data address, strides) of the device store
        self._store_refs = (None, None)   # weak references to the (data_search, pred_categories) objects it was built from
        self._cat_sig = None              # buffer signature of the data_navigation['category'] column last seen by search_single
        self._nav_side_effect = (None, None)
        self.store_from_cache = False
        self.store_builds = 0             # device stores built or loaded so far (tests assert reuse with it)

    # ---- lifetime ----
    def close(self):
        if getattr(self, "_h", None) and self._h.value:
            self._lib.lmi_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_option(self, name, value):
        check(self._lib.lmi_set_option(self._h, name.encode(), int(value)))

    # ---- model ----
    def set_model(self, weights, biases):
        """Load the nn.Linear stack (weights[l] = [out,in] fp32).  Sets ``self.model``."""
        w = [np.ascontiguousarray(np.asarray(x, dtype=np.float32)) for x in weights]
        b = [np.ascontiguousarray(np.asarray(x, dtype=np.float32)) for x in biases]
        if w[0].shape[1] != self.d_nav:
            raise ValueError("first layer input width must equal d_nav")
        for i in range(1, len(w)):
            if w[i].shape[1] != w[i - 1].shape[0]:
                raise ValueError("layer shapes do not chain")
        n = len(w)
        dims = (C.c_int32 * n)(*[x.shape[0] for x in w])
        wp = (C.c_void_p * n)(*[x.ctypes.data for x in w])
        bp = (C.c_void_p * n)(*[x.ctypes.data for x in b])
        check(self._lib.lmi_set_mlp(self._h, n, dims, wp, bp))
        self.model = GpuNavigator(self)
        return self

    @classmethod
    def from_reference(cls, li, n_categories=None, d_nav=None, **kw):
        """Build from an (unmodified) reference ``li.LearnedIndex`` after ``build`` / ``pickle.load``."""
        w, b = extract_mlp(li)
        obj = cls(n_categories or w[-1].shape[0], d_nav=d_nav or w[0].shape[1], **kw)
        return obj.set_model(w, b)

    # ---- store ----
    def _content_key(self, rows, id_base, pc):
        """32-byte fingerprint of what the device store is a function of: ALL of pred_categories (8 bytes x N is
        cheap next to the upload), the shapes, id_base, the shard, and 66 sampled rows of the data.  Names the
        on-disk cache file and is stored inside it."""
        h = hashlib.blake2b(digest_size=32)
        h.update(memoryview(pc).cast("B"))
        n = rows.shape[0]
        h.update(struct.pack("<qqqiii", n, rows.shape[1], id_base, self.rank, self.world, self.shard_policy))
        h.update(str(rows.dtype).encode())
        if n:
            pick = np.unique(np.concatenate([np.arange(0, n, max(1, n // 64)), [n - 1]]))
            h.update(np.ascontiguousarray(rows[pick]).tobytes())
        return h.digest()

    def invalidate(self):
        """Forget the cached store: the next prepare()/search() rebuilds it (call after editing the arrays in place)."""
        self._store_key = None
        self._store_refs = (None, None)
        self._cat_sig = None

    def prepare(self, data_search, pred_categories, force=False, cache_dir=None):
        """Build (or reuse) the device bucket store for (data_search, pred_categories).  The reference regroups and
        regathers on every rank of every search (LearnedIndex.py:143,152); here it happens once per dataset.

        Reuse rules (nothing below converts or copies ``data_search``): (1) the very same live objects as last time ->
        reused at once; (2) otherwise a full hash of ``pred_categories`` plus the data's buffer address, layout and 66
        sampled rows are compared with the store's -> reused if equal (e.g. ``df.values`` temporaries); (3) else
        rebuilt.  In-place edits of the same objects are not seen: pass ``force=True`` or call ``invalidate()``.
        ``cache_dir``: keep / look up the finished device layout on disk (``lmi_store_save`` / ``lmi_store_load``),
        file name = the content key."""
        if not force and self._store_key is not None and _alive(self._store_refs[0], data_search) and \
                _alive(self._store_refs[1], pred_categories):
            return self
        pc = np.ascontiguousarray(np.asarray(pred_categories, dtype=np.int64))
        rows, id_base = _view_rows(data_search)
        if rows.ndim != 2 or rows.shape[1] != self.d:
            raise ValueError(f"data_search must be [N, {self.d}]")
        if pc.shape[0] != rows.shape[0]:
            raise ValueError("pred_categories length != number of objects")
        ckey = self._content_key(rows, id_base, pc)
        key = (ckey, rows.__array_interface__["data"][0], rows.strides)
        if not force and key == self._store_key:
            self._store_refs = (_weak(data_search), _weak(pred_categories))
            return self
        if rows.dtype != np.float16:
            rows = rows.astype(np.float16)            # clip768v2 'emb' is float16 (search.py:79-81)
        es = rows.dtype.itemsize
        if not ((rows.strides[1] == es and rows.strides[0] >= self.d * es and rows.strides[0] % es == 0) or
                (rows.strides[0] == es and rows.strides[1] >= rows.shape[0] * es and rows.strides[1] % es == 0)):
            rows = np.ascontiguousarray(rows)         # exotic strides only; frames and arrays never get here
        path = None
        if cache_dir is not None:
            os.makedirs(cache_dir, exist_ok=True)
            path = os.path.join(cache_dir, f"lmi_store_{ckey.hex()[:32]}.bin")
        loaded = False
        if path is not None and os.path.exists(path):
            loaded = self._lib.lmi_store_load(self._h, path.encode(), ckey) == 0     # any mismatch -> rebuild below
        if not loaded:
            n = rows.shape[0]
            check(self._lib.lmi_store_begin(self._h, _ptr(pc), n, id_base, 0))
            check(self._lib.lmi_store_append_strided(self._h, C.c_void_p(rows.ctypes.data), 0, n,
                                                     rows.strides[0] // es, rows.strides[1] // es, 0))
            check(self._lib.lmi_store_end(self._h))
            if path is not None:
                check(self._lib.lmi_store_save(self._h, path.encode(), ckey))
        self.store_from_cache = loaded
        self.store_builds += 1
        self._store_key = key
        self._store_refs = (_weak(data_search), _weak(pred_categories))
        self._cat_sig = None
        return self

    def prepare_device(self, data_f16_t, pred_categories_t, id_base=1, chunk_rows=1 << 20):
        """Same from torch CUDA tensors already in HBM (fp16 [N,d], int64 [N])."""
        n = int(data_f16_t.shape[0])
        check(self._lib.lmi_store_begin(self._h, C.c_void_p(pred_categories_t.data_ptr()), n, id_base, 1))
        import torch
        torch.cuda.current_stream().synchronize()
        for lo in range(0, n, chunk_rows):
            cnt = min(chunk_rows, n - lo)
            check(self._lib.lmi_store_append(self._h, C.c_void_p(data_f16_t[lo:lo + cnt].data_ptr()), lo, cnt, 1))
        check(self._lib.lmi_store_end(self._h))
        self._store_key = ("device", n)
        self._store_refs = (None, None)
        self._cat_sig = None
        return self

    def bucket_ptr(self):
        out = np.empty(self.n_categories + 1, dtype=np.int64)
        check(self._lib.lmi_get_bucket_ptr(self._h, _ptr(out)))
        return out

    # ---- search (reference: LearnedIndex.py:22-101) ----
    def search(self, data_navigation, queries_navigation, data_search, queries_search, pred_categories,
               n_buckets=1, k=10, use_threshold=False):
        assert self.model is not None, 'Model is not trained, call `build` first.'      # LearnedIndex.py:60
        if data_navigation is not None and hasattr(data_navigation, "columns"):
            # LearnedIndex.py:67 mutates the caller's frame on every call (an 80 MB column copy at 10M); skipped when
            # the same live frame already received the same live pred_categories object from the previous call
            same = _alive(self._nav_side_effect[0], data_navigation) and _alive(self._nav_side_effect[1], pred_categories) \
                and "category" in data_navigation.columns
            if not same:
                data_navigation['category'] = pred_categories
                self._nav_side_effect = (_weak(data_navigation), _weak(pred_categories))
        self.prepare(data_search, pred_categories)
        return self._search_prepared(queries_navigation, queries_search, n_buckets, k)

    def _search_prepared(self, queries_navigation, queries_search, n_buckets, k):
        """The search proper on the store prepared last (host buffers in and out through ``lmi_search``)."""
        qn = np.ascontiguousarray(np.asarray(queries_navigation, dtype=np.float32))      # model.py:88
        qs = np.ascontiguousarray(np.asarray(queries_search, dtype=np.float16))
        nq = qs.shape[0]
        if qn.shape != (nq, self.d_nav) or qs.shape != (nq, self.d):
            raise ValueError("query arrays have the wrong shape")
        dists = np.empty((nq, k), dtype=np.float64)
        anns = np.empty((nq, k), dtype=np.uint32)
        check(self._lib.lmi_search(self._h, _ptr(qn), _ptr(qs.view(np.uint16)), nq, int(n_buckets), int(k),
                                   _ptr(dists), _ptr(anns)))
        assert anns.shape == dists.shape == (nq, k)                                      # LearnedIndex.py:99
        return dists, anns

    # ---- search_single (reference: LearnedIndex.py:103-195) ----
    def search_single(self, data_navigation, data_search, queries_search, pred_categories, k=10,
                      threshold_dist=None):
        """``pred_categories`` here is the category of every *query* at one rank (as in the reference).
        The object->bucket assignment is ``data_navigation['category']`` (the reference regroups by it on every call,
        LearnedIndex.py:143): the store is reused when that column is the one it was built from (same frame, same
        column buffer; else its full hash is compared) and rebuilt when it changed.  Without the column (``None``
        frame) a store prepared earlier is used."""
        cat_col = None
        if data_navigation is not None and "category" in getattr(data_navigation, "columns", []):
            cat_col = data_navigation["category"].to_numpy()
        if cat_col is None:
            if self._store_key is None:
                raise AssertionError("no bucket store: call prepare() or pass data_navigation with a 'category' column")
        else:
            sig = (cat_col.__array_interface__["data"][0], cat_col.shape, cat_col.strides, str(cat_col.dtype))
            if not (self._store_key is not None and sig == self._cat_sig and _alive(self._store_refs[0], data_search)):
                self._store_refs = (None, None)                 # force the content comparison (full hash of the column)
                self.prepare(data_search, cat_col)
                self._cat_sig = sig
        qs = np.ascontiguousarray(np.asarray(queries_search, dtype=np.float16))
        nq = qs.shape[0]
        cat = np.ascontiguousarray(np.asarray(pred_categories, dtype=np.int64))
        thr = None if threshold_dist is None else np.ascontiguousarray(np.asarray(threshold_dist, dtype=np.float64))
        dists = np.empty((nq, k), dtype=np.float64)
        nns = np.empty((nq, k), dtype=np.uint32)
        check(self._lib.lmi_search_rank(self._h, _ptr(qs.view(np.uint16)), nq, _ptr(cat), int(k), _ptr(thr),
                                        _ptr(dists), _ptr(nns)))
        return dists, nns

    # ---- device-resident variants (benchmarks, multi-GPU) ----
    def search_device(self, q_nav_t, q_f16_t, n_buckets, k, out=None):
        import torch
        nq = int(q_f16_t.shape[0])
        if out is None:
            out = (torch.empty(nq, k, dtype=torch.float64, device=q_f16_t.device),
                   torch.empty(nq, k, dtype=torch.int32, device=q_f16_t.device))
        st = torch.cuda.current_stream(q_f16_t.device).cuda_stream
        check(self._lib.lmi_search_device(self._h, C.c_void_p(q_nav_t.data_ptr()), C.c_void_p(q_f16_t.data_ptr()), nq,
                                          int(n_buckets), int(k), C.c_void_p(out[0].data_ptr()),
                                          C.c_void_p(out[1].data_ptr()), C.c_void_p(st)))
        return out

    def search_partial_device(self, q_nav_t, q_f16_t, n_buckets, k, partial_t):
        import torch
        nq = int(q_f16_t.shape[0])
        st = torch.cuda.current_stream(q_f16_t.device).cuda_stream
        check(self._lib.lmi_search_partial_device(self._h, C.c_void_p(q_nav_t.data_ptr()), C.c_void_p(q_f16_t.data_ptr()),
                                                  nq, int(n_buckets), int(k), C.c_void_p(partial_t.data_ptr()), C.c_void_p(st)))

    def navigate_device(self, q_nav_t, n_top, out_classes_t):
        """top-n_top bucket ids (int32 [rows, n_top], device) of device-resident navigation vectors."""
        import torch
        st = torch.cuda.current_stream(q_nav_t.device).cuda_stream
        check(self._lib.lmi_navigate_device(self._h, C.c_void_p(q_nav_t.data_ptr()), int(q_nav_t.shape[0]), int(n_top),
                                            C.c_void_p(out_classes_t.data_ptr()), C.c_void_p(st)))

    def search_partial_classes_device(self, classes_t, q_f16_t, n_buckets, k, partial_t):
        """search_partial_device with caller-supplied bucket lists (int32 [nq, n_buckets], device)."""
        import torch
        nq = int(q_f16_t.shape[0])
        st = torch.cuda.current_stream(q_f16_t.device).cuda_stream
        check(self._lib.lmi_search_partial_classes_device(self._h, C.c_void_p(classes_t.data_ptr()), C.c_void_p(q_f16_t.data_ptr()),
                                                          nq, int(n_buckets), int(k), C.c_void_p(partial_t.data_ptr()), C.c_void_p(st)))

    def merge_partials_device(self, gathered_t, world, nq, k, out):
        import torch
        st = torch.cuda.current_stream(gathered_t.device).cuda_stream
        check(self._lib.lmi_merge_partials_device(self._h, C.c_void_p(gathered_t.data_ptr()), int(world), int(nq), int(k),
                                                  C.c_void_p(out[0].data_ptr()), C.c_void_p(out[1].data_ptr()), C.c_void_p(st)))
        return out

    def partial_bytes(self, nq, k):
        return int(self._lib.lmi_partial_bytes(int(nq), int(k)))

    def stats(self):
        s = LmiStats()
        check(self._lib.lmi_get_stats(self._h, C.byref(s)))
        d = s.as_dict()
        d["rescore_kernel"] = self._lib.lmi_rescore_kernel_name(self._h).decode()
        return d


class GpuBaseline:
    """Exhaustive search with the rescoring kernels: mirror of ``li.Baseline`` (search/li/Baseline.py:7-25;
    SURVEY.md 8(f) row N4).  ``search(queries, data, k)`` -> ``(dists f64 [nq,k], nns [nq,k] 1-based, seconds)``
    = the k smallest ``1 - cos`` over ALL objects.  Internally a one-bucket index (every object in category
    0, every query routed to it), so the same tcgen05 kernel, float64 refine and certificate apply; equal
    distances resolve to the lower id (the reference's ``argsort`` order among exact ties is unspecified).
    ``bench.py``'s ground truth deliberately does NOT use this class (it must stay independent of K3)."""

    def __init__(self, device=0, rescore=RESCORE_TC):
        self.device, self.rescore = device, rescore
        self._li, self._key = None, None

    def build(self, data):
        """Baseline.build is a no-op that returns its duration (Baseline.py:22-25); here it uploads."""
        import time
        s = time.time()
        rows, _ = _as_rows(data)
        key = (rows.shape, hashlib.blake2b(rows.tobytes(), digest_size=16).digest())     # (documented for < 1M objects, Baseline.py:8-10)
        if key != self._key:
            if self._li is not None:
                self._li.close()
            self._li = GpuLearnedIndex(1, d=rows.shape[1], device=self.device, rescore=self.rescore)
            # Baseline.search numbers the rows 1..N regardless of the DataFrame's labels (Baseline.py:18)
            self._li.prepare(np.asarray(rows), np.zeros(rows.shape[0], dtype=np.int64))
            self._key = key
        return time.time() - s

    def search(self, queries, data, k=10):
        import time
        s = time.time()
        self.build(data)
        qs = np.asarray(queries)
        dists, nns = self._li.search_single(None, None, qs, np.zeros(qs.shape[0], dtype=np.int64), k=k)
        return dists, nns, time.time() - s
