"""Multi-GPU layer of the query path: one process per GPU, buckets (or bucket stripes) sharded
across ranks, ONE exchange step per search (an all-gather of every rank's top-k), then a merge.

The reference has nothing distributed; this is new.  It is exact because
buckets are disjoint object sets (``pred_categories`` assigns one category per object,
search/li/LearnedIndex.py:240) and the answer is a top-k over their union, so the top-k of the
per-rank top-k lists is the global top-k.  The merge key (float64 distance, bucket rank, id) makes
the result independent of the number of ranks.

Host-side pieces (numpy, no GPU needed -- these are what the world_size-2 ``gloo`` tests run):
  * ``lpt_owners``      -- the bucket -> rank map, identical to the one liblmi_b200.so computes in
                           ``lmi_store_begin`` (csrc/lmi_api_store.inl: longest-processing-time on n_b^2)
  * ``shard_keep_mask`` -- which objects a rank holds under each policy (whole buckets, stripes of every
                           bucket, or caller-partitioned row ranges -- ``SHARD_RANGES``, for datasets that
                           never exist in one piece)
  * ``merge_partials``  -- numpy mirror of ``merge_partials_kernel`` (csrc/finalize.cuh)
Device piece:
  * ``DistributedSearch`` -- search_partial -> ``dist.all_gather_into_tensor`` (NCCL) -> merge kernel
"""
from __future__ import annotations

import numpy as np

SHARD_BUCKETS, SHARD_STRIPES = 0, 1
# Caller-partitioned objects: rank r is handed rows [N*r/world, N*(r+1)/world) only and builds an ordinary
# (unsharded) index over them with id_base = first row + 1.  This is how a dataset that does not fit one
# host or one GPU is loaded (100M x 768 fp16 = 154 GB: 19.2 GB per GPU); the exchange and merge are the same.
SHARD_RANGES = 2

# layout of one exchanged entry (csrc/common.cuh: struct Partial, 16 bytes)
PARTIAL_DTYPE = np.dtype([("dist", "<f8"), ("id", "<u4"), ("rank", "<u4")])
EMPTY_DIST = 10_000.0            # LearnedIndex.py:138
EMPTY_RANK = 0xFFFFFFFF


def lpt_owners(sizes, world):
    """Bucket -> rank, longest-processing-time first on weight n_b^2 (expected work of a bucket
    ~ n_b * m_b and queries follow the data distribution).  Stable for equal sizes."""
    sizes = np.asarray(sizes, dtype=np.int64)
    order = np.argsort(-sizes, kind="stable")
    load = np.zeros(world, dtype=np.float64)
    owner = np.zeros(sizes.shape[0], dtype=np.int32)
    for c in order:
        best = int(np.argmin(load))          # first minimum, like the C++ loop
        owner[c] = best
        load[best] += float(sizes[c]) * float(sizes[c])
    return owner


def shard_keep_mask(pred_categories, n_categories, rank, world, policy=SHARD_BUCKETS):
    """Boolean mask over objects: True where this rank holds the object.
    buckets: whole buckets by ``lpt_owners``; stripes: rows [n*rank/world, n*(rank+1)/world) of
    every bucket, in ascending object order (csrc/store.cuh: store_layout_kernel)."""
    pc = np.asarray(pred_categories, dtype=np.int64)
    if world <= 1:
        return np.ones(pc.shape[0], dtype=bool)
    sizes = np.bincount(pc, minlength=n_categories).astype(np.int64)
    if policy == SHARD_BUCKETS:
        return lpt_owners(sizes, world)[pc] == rank
    if policy == SHARD_RANGES:
        lo, hi = range_of_rank(pc.shape[0], rank, world)
        i = np.arange(pc.shape[0])
        return (i >= lo) & (i < hi)
    # rank of every object inside its bucket (stable order = ascending object id)
    order = np.argsort(pc, kind="stable")
    start = np.concatenate([[0], np.cumsum(sizes)[:-1]])
    in_bucket = np.empty(pc.shape[0], dtype=np.int64)
    in_bucket[order] = np.arange(pc.shape[0]) - start[pc[order]]
    n = sizes[pc]
    lo, hi = n * rank // world, n * (rank + 1) // world
    return (in_bucket >= lo) & (in_bucket < hi)


def range_of_rank(n, rank, world):
    """Rows [lo, hi) of an N-row dataset that rank `rank` loads under SHARD_RANGES."""
    return n * rank // world, n * (rank + 1) // world


def query_slice_of_rank(nq, rank, world):
    """Navigation split: rank r navigates query rows [lo, hi) = [r*per, (r+1)*per) clipped to nq, per = ceil(nq/world);
    slices are padded to `per` rows so that one all-gather of [per, n_buckets] int32 blocks rebuilds the whole list
    (rows >= nq are padding at the very end).  -> (per, lo, hi)"""
    per = (nq + world - 1) // world
    return per, min(nq, rank * per), min(nq, (rank + 1) * per)


def pack_partials(dists, ids, ranks):
    """(f64 [nq,k], u32 [nq,k], u32 [nq,k]) -> structured array [nq,k] in the exchange layout."""
    out = np.empty(dists.shape, dtype=PARTIAL_DTYPE)
    out["dist"], out["id"], out["rank"] = dists, ids, ranks
    return out


def merge_partials(gathered, k):
    """gathered: structured [world, nq, k] (rank-major, as the all-gather delivers it).
    Returns (dists f64 [nq,k], ids u32 [nq,k]): the k smallest by (dist, bucket rank, id, source
    position) -- the total order ``key_less`` of csrc/finalize.cuh."""
    world, nq, kk = gathered.shape
    flat = np.transpose(gathered, (1, 0, 2)).reshape(nq, world * kk)
    src = np.broadcast_to(np.arange(world * kk), flat.shape)
    order = np.lexsort((src, flat["id"], flat["rank"], flat["dist"]), axis=1)[:, :k]
    sel = np.take_along_axis(flat, order, axis=1)
    return sel["dist"].astype(np.float64), sel["id"].astype(np.uint32)


class DistributedSearch:
    """One search step on ``world`` GPUs: every rank navigates all queries, rescoring covers the
    rank's own buckets, one ``all_gather_into_tensor`` of [nq, k] x 16 B, merge on every rank."""

    def __init__(self, index, group=None, slice_navigation=True):
        """slice_navigation: every rank navigates only rows [rank*per, (rank+1)*per) of the batch and the bucket
        lists (nq x n_buckets int32: 160 KB at 10k x 4) are all-gathered, instead of every rank running the MLP
        over all queries -- navigation is replicated work that does not shrink with the number of GPUs (7.6 % of an
        8-GPU step in round 1).  Same kernel on the same rows, so the lists, and the results, are bit-identical."""
        import torch.distributed as dist
        self.li, self.dist, self.group = index, dist, group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.slice_navigation = slice_navigation
        self._bufs = None
        self._cls = None
        # search_device_pipelined: exchange stream, its own communicator, two sets of exchange buffers
        self._xs = None
        self._xgroup = None
        self._ring = None
        self._ring_done = [None, None]
        self._n_pipelined = 0
        # navigation look-ahead (search_device_pipelined(next_q_nav=...)): two sets of list buffers, the prepared batch
        self._la_bufs = None
        self._la_read = [None, None]
        self._la_n = 0
        self._la_pending = None

    def search_device(self, q_nav_t, q_f16_t, n_buckets, k, out=None):
        import torch
        nq = int(q_f16_t.shape[0])
        dev = q_f16_t.device
        if out is None:
            out = (torch.empty(nq, k, dtype=torch.float64, device=dev), torch.empty(nq, k, dtype=torch.int32, device=dev))
        if self.world == 1:
            return self.li.search_device(q_nav_t, q_f16_t, n_buckets, k, out)
        nbytes = self.li.partial_bytes(nq, k)
        if self._bufs is None or self._bufs[0].numel() != nbytes:
            self._bufs = (torch.empty(nbytes, dtype=torch.uint8, device=dev),
                          torch.empty(self.world * nbytes, dtype=torch.uint8, device=dev))
        part, gathered = self._bufs
        if self.slice_navigation:
            every = self.navigate_sliced(q_nav_t, n_buckets)
            self.li.search_partial_classes_device(every, q_f16_t, n_buckets, k, part)   # (rows >= nq are padding, never read)
        else:
            self.li.search_partial_device(q_nav_t, q_f16_t, n_buckets, k, part)
        self.dist.all_gather_into_tensor(gathered, part, group=self.group)      # the exchange step of the search proper
        return self.li.merge_partials_device(gathered, self.world, nq, k, out)

    def search_device_pipelined(self, q_nav_t, q_f16_t, n_buckets, k, out, next_q_nav=None):
        """Throughput form of ``search_device`` for a stream of batches: the exchange step of batch i (all-gather of the
        partial results + merge kernel) runs on a side stream while the caller's stream goes on to batch i + 1, whose
        navigation, bucket-list all-gather and planning use the GPU only lightly.  Per-rank results, merge key and
        therefore the output are those of ``search_device``; only the stream the merge runs on differs.

        Returns a ``torch.cuda.Event`` recorded behind the merge: ``out`` may be read once it has completed
        (``event.wait()`` makes the current stream wait for it, ``event.synchronize()`` the host).  The partial and
        gathered buffers are double-buffered; a third call waits for the first one's exchange.  The exchange uses its
        own NCCL communicator (created collectively on the first call) so that it does not queue behind, or ahead of,
        the next batch's bucket-list all-gather on one NCCL stream.  Every rank must issue the same calls in the same
        order, as with any collective.

        ``next_q_nav`` (navigation look-ahead, sliced navigation only): the navigation vectors of the batch the NEXT call
        will search.  This rank's slice of them goes through K1 now, ahead of this batch's rescoring, and the all-gather
        of the bucket lists is queued on the exchange stream, where it runs beside this batch's finalize -- so the next
        call finds its bucket lists ready and no collective is left on the search stream.  The next call must pass the
        same tensor (same storage and shape) as its ``q_nav_t``; otherwise the prepared lists are ignored."""
        import torch
        nq = int(q_f16_t.shape[0])
        dev = q_f16_t.device
        if self.world == 1:
            self.li.search_device(q_nav_t, q_f16_t, n_buckets, k, out)
            ev = torch.cuda.Event()
            ev.record(torch.cuda.current_stream(dev))
            return ev
        if self._xs is None:
            self._xs = torch.cuda.Stream(device=dev)
            ranks = None if self.group is None else self.dist.get_process_group_ranks(self.group)
            self._xgroup = self.dist.new_group(ranks=ranks)
        nbytes = self.li.partial_bytes(nq, k)
        if self._ring is None or self._ring[0][0].numel() != nbytes:
            self.flush_pipelined()
            self._ring = [(torch.empty(nbytes, dtype=torch.uint8, device=dev),
                           torch.empty(self.world * nbytes, dtype=torch.uint8, device=dev)) for _ in range(2)]
            self._ring_done = [None, None]
        slot = self._n_pipelined & 1
        self._n_pipelined += 1
        part, gathered = self._ring[slot]
        main = torch.cuda.current_stream(dev)
        if self._ring_done[slot] is not None:
            main.wait_event(self._ring_done[slot])           # the exchange two batches back has read `part`
        if self.slice_navigation:
            key = (q_nav_t.data_ptr(), tuple(q_nav_t.shape), int(n_buckets))
            la = self._la_pending
            if la is not None and la[0] == key:
                every = la[1]
                main.wait_event(la[2])                       # bucket lists gathered on the exchange stream a batch ago
            else:
                every = self.navigate_sliced(q_nav_t, n_buckets)
            self._la_pending = None
            if next_q_nav is not None:
                self._la_pending = self._navigate_ahead(next_q_nav, n_buckets, main)
            self.li.search_partial_classes_device(every, q_f16_t, n_buckets, k, part)
        else:
            self.li.search_partial_device(q_nav_t, q_f16_t, n_buckets, k, part)
        ready = torch.cuda.Event()
        ready.record(main)
        done = torch.cuda.Event()
        with torch.cuda.stream(self._xs):
            self._xs.wait_event(ready)
            self.dist.all_gather_into_tensor(gathered, part, group=self._xgroup)
            self.li.merge_partials_device(gathered, self.world, nq, k, out)
            done.record(self._xs)
        self._ring_done[slot] = done
        return done

    def _navigate_ahead(self, q_nav_t, n_buckets, main):
        """K1 on this rank's slice of a FUTURE batch on the search stream, all-gather of the lists on the exchange
        stream.  -> (key, gathered lists, event recorded behind the all-gather)"""
        import torch
        nq, dev, W = int(q_nav_t.shape[0]), q_nav_t.device, self.world
        r = self.dist.get_rank(self.group)
        per, lo, hi = query_slice_of_rank(nq, r, W)
        if self._la_bufs is None or self._la_bufs[0][0].shape != (per, n_buckets):
            self._la_bufs = [(torch.full((per, n_buckets), -1, dtype=torch.int32, device=dev),
                              torch.empty(W * per, n_buckets, dtype=torch.int32, device=dev)) for _ in range(2)]
            self._la_read = [None, None]
        slot = self._la_n & 1
        self._la_n += 1
        mine, every = self._la_bufs[slot]
        if self._la_read[slot] is not None:
            main.wait_event(self._la_read[slot])             # the all-gather that last read `mine` has finished
        if hi > lo:
            self.li.navigate_device(q_nav_t[lo:hi], n_buckets, mine)
        nav_done = torch.cuda.Event()
        nav_done.record(main)
        gathered = torch.cuda.Event()
        with torch.cuda.stream(self._xs):
            self._xs.wait_event(nav_done)
            self.dist.all_gather_into_tensor(every, mine, group=self._xgroup)
            gathered.record(self._xs)
        self._la_read[slot] = gathered
        return ((q_nav_t.data_ptr(), tuple(q_nav_t.shape), int(n_buckets)), every, gathered)

    def flush_pipelined(self):
        """The current stream waits for every exchange issued by ``search_device_pipelined``."""
        for ev in self._ring_done:
            if ev is not None:
                ev.wait()

    def navigate_sliced(self, q_nav_t, n_buckets):
        """This rank navigates its slice of the batch; ONE all-gather of [per, n_buckets] int32 blocks rebuilds the
        bucket lists of all queries (int32 [world * per, n_buckets], rows >= nq are -1 padding)."""
        import torch
        nq, dev, W = int(q_nav_t.shape[0]), q_nav_t.device, self.world
        r = self.dist.get_rank(self.group)
        per, lo, hi = query_slice_of_rank(nq, r, W)
        if self._cls is None or self._cls[0].shape != (per, n_buckets):
            self._cls = (torch.full((per, n_buckets), -1, dtype=torch.int32, device=dev),
                         torch.empty(W * per, n_buckets, dtype=torch.int32, device=dev))
        mine, every = self._cls
        if hi > lo:
            self.li.navigate_device(q_nav_t[lo:hi], n_buckets, mine)
        self.dist.all_gather_into_tensor(every, mine, group=self.group)         # bucket lists: nq x n_buckets x 4 bytes
        return every

    def search_host(self, q_nav_host, q_f16_host, n_buckets, k, out_host=None):
        """End-to-end step from (pinned) host tensors that every rank holds: each rank uploads only ITS slice
        of the queries over PCIe and the ranks all-gather the slices over NVLink (19.2 MB at 10k queries: tens
        of microseconds instead of a full upload per rank), then the sharded search, then the merged result is
        copied back to the host.  Returns (dists f64 [nq,k], ids i32 [nq,k]) as pinned host tensors."""
        import torch
        nq = int(q_f16_host.shape[0])
        dev = torch.device("cuda", self.li.device)
        W, r = self.world, (self.dist.get_rank(self.group) if self.world > 1 else 0)
        per = (nq + W - 1) // W                                  # rows per rank (last slice zero-padded)
        key = (nq, int(q_nav_host.shape[1]), int(q_f16_host.shape[1]), k)
        if getattr(self, "_hbuf_key", None) != key:
            self._hbuf_key = key
            self._qn_all = torch.empty(W * per, q_nav_host.shape[1], dtype=torch.float32, device=dev)
            self._qs_all = torch.empty(W * per, q_f16_host.shape[1], dtype=torch.float16, device=dev)
            self._qn_mine = torch.zeros(per, q_nav_host.shape[1], dtype=torch.float32, device=dev)
            self._qs_mine = torch.zeros(per, q_f16_host.shape[1], dtype=torch.float16, device=dev)
            self._out_dev = (torch.empty(nq, k, dtype=torch.float64, device=dev), torch.empty(nq, k, dtype=torch.int32, device=dev))
            self._out_host = (torch.empty(nq, k, dtype=torch.float64).pin_memory(), torch.empty(nq, k, dtype=torch.int32).pin_memory())
        lo, hi = min(nq, r * per), min(nq, (r + 1) * per)
        if W == 1:
            qn_all, qs_all = self._qn_all[:nq], self._qs_all[:nq]
            qn_all.copy_(q_nav_host, non_blocking=True); qs_all.copy_(q_f16_host, non_blocking=True)
        else:
            if hi > lo:
                self._qn_mine[:hi - lo].copy_(q_nav_host[lo:hi], non_blocking=True)
                self._qs_mine[:hi - lo].copy_(q_f16_host[lo:hi], non_blocking=True)
            self.dist.all_gather_into_tensor(self._qn_all, self._qn_mine, group=self.group)
            self.dist.all_gather_into_tensor(self._qs_all, self._qs_mine, group=self.group)
            qn_all, qs_all = self._qn_all[:nq], self._qs_all[:nq]
        self.search_device(qn_all, qs_all, n_buckets, k, self._out_dev)
        oh = out_host if out_host is not None else self._out_host
        oh[0].copy_(self._out_dev[0], non_blocking=True); oh[1].copy_(self._out_dev[1], non_blocking=True)
        torch.cuda.current_stream(dev).synchronize()
        return oh
