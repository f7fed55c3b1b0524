"""N>1 host logic on CPU: world_size-2 (and 3) ``gloo`` runs of the sharded search protocol.

Every rank scores only the objects it holds (``shard_keep_mask``), the per-rank top-k lists are
exchanged with ONE all-gather in the exchange layout (``PARTIAL_DTYPE``), and ``merge_partials``
must reproduce the single-rank result bit for bit -- ids and float64 distances -- under both shard
policies.  The per-rank scorer here is the CPU oracle (tests may use it as the checker); on the
GPU box the same protocol runs through lmi_search_partial_device / NCCL / merge_partials_kernel
(tests/test_gpu_seeded.py::test_sharded_merge_equals_single, bench.py --gpus N).
"""
import os
import socket
import sys

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
for p in (ROOT, HERE):
    if p not in sys.path:
        sys.path.insert(0, p)


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _partial_for_rank(g, nb, rank, world, policy):
    """Top-k of the objects held by `rank`, with the bucket rank of every hit (the merge key)."""
    from oracle import lmi_oracle as O
    from lmi_b200 import shard as S
    pc = g["pred_categories"].astype(np.int64)
    keep = S.shard_keep_mask(pc, g["n_categories"], rank, world, policy)
    pc_local = np.where(keep, pc, -1)                      # objects held elsewhere match no category
    k = g["k"]
    d, a = O.search_intended(g["classes"], g["data_f16"], g["queries_f16"], pc_local, nb, k)
    cls = g["classes"][:, :nb]
    rk = np.full(a.shape, S.EMPTY_RANK, dtype=np.uint32)
    filled = a > 0
    cat_of_hit = np.where(filled, pc[np.maximum(a.astype(np.int64) - 1, 0)], -2)
    for r in range(nb):
        rk[filled & (cat_of_hit == cls[:, r:r + 1])] = r
    assert (rk[filled] != S.EMPTY_RANK).all()
    return S.pack_partials(d, a, rk), keep


def _worker(rank, world, port, tag, policy, out_dir):
    import torch
    import torch.distributed as dist
    from conftest import load_golden
    from lmi_b200 import shard as S
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    try:
        g = load_golden(tag)
        results = {}
        from oracle import lmi_oracle as O
        for nb in g["bucket_list"]:
            # navigation split (DistributedSearch.slice_navigation): every rank navigates its slice of the queries,
            # ONE all-gather of [per, nb] int32 blocks rebuilds the bucket lists of the whole batch
            nq = g["queries_nav"].shape[0]
            per, lo, hi = S.query_slice_of_rank(nq, rank, world)
            mine = np.full((per, nb), -1, dtype=np.int32)
            if hi > lo:
                mine[:hi - lo] = O.predict_proba(g["queries_nav"][lo:hi], g["weights"], g["biases"])[1][:, :nb]
            every = torch.empty(world * per, nb, dtype=torch.int32)
            dist.all_gather_into_tensor(every, torch.from_numpy(mine))
            every = every.numpy()
            assert (every[:nq] == g["classes"][:, :nb]).all() and (every[nq:] == -1).all()
            part, keep = _partial_for_rank(g, nb, rank, world, policy)
            # every object is held by exactly one rank
            cnt = torch.from_numpy(keep.astype(np.int32))
            dist.all_reduce(cnt)
            assert bool((cnt == 1).all())
            # the one exchange step: all-gather of [nq, k] x 16 B, rank-major
            mine = torch.from_numpy(part.view(np.uint8).reshape(-1).copy())
            gathered = torch.empty(world * mine.numel(), dtype=torch.uint8)
            dist.all_gather_into_tensor(gathered, mine)
            gathered = gathered.numpy().view(S.PARTIAL_DTYPE).reshape(world, *part.shape)
            results[nb] = S.merge_partials(gathered, g["k"])
        np.savez(os.path.join(out_dir, f"rank{rank}.npz"),
                 **{f"d{nb}": v[0] for nb, v in results.items()}, **{f"a{nb}": v[1] for nb, v in results.items()})
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("policy", [0, 1, 2], ids=["buckets", "stripes", "ranges"])
@pytest.mark.parametrize("world", [2, 3])
def test_sharded_protocol_matches_single_rank(tmp_path, world, policy):
    import torch.multiprocessing as mp
    from oracle import lmi_oracle as O
    from conftest import load_golden
    tag = "ref_mlp_n2400_c24"
    mp.spawn(_worker, args=(world, _free_port(), tag, policy, str(tmp_path)), nprocs=world, join=True)
    g = load_golden(tag)
    for nb in g["bucket_list"]:
        d1, a1 = O.search_intended(g["classes"], g["data_f16"], g["queries_f16"], g["pred_categories"], nb, g["k"])
        for r in range(world):
            z = np.load(os.path.join(str(tmp_path), f"rank{r}.npz"))
            assert np.array_equal(z[f"a{nb}"], a1), f"ids differ on rank {r}, nb={nb}"
            if policy == 0:
                assert np.array_equal(z[f"d{nb}"], d1), f"distances differ on rank {r}, nb={nb}"
            else:
                # stripes / ranges change the shape of the checker's BLAS products (last-ulp differences in
                # numpy, not in the protocol; the CUDA path scores every pair identically)
                np.testing.assert_allclose(z[f"d{nb}"], d1, rtol=0, atol=1e-12)


def test_lpt_owners_balances_and_is_deterministic():
    from lmi_b200 import shard as S
    rng = np.random.default_rng(0)
    sizes = rng.integers(0, 5000, size=122)
    for world in (2, 4, 8):
        o1, o2 = S.lpt_owners(sizes, world), S.lpt_owners(sizes, world)
        assert np.array_equal(o1, o2) and o1.min() >= 0 and o1.max() < world
        load = np.array([(sizes[o1 == r].astype(np.float64) ** 2).sum() for r in range(world)])
        assert load.max() <= load.mean() * 1.25            # LPT: well within 4/3 of the mean here


def test_stripes_partition_every_bucket():
    from lmi_b200 import shard as S
    rng = np.random.default_rng(1)
    pc = rng.integers(0, 16, size=5000)
    for world in (2, 3, 8):
        masks = np.stack([S.shard_keep_mask(pc, 16, r, world, S.SHARD_STRIPES) for r in range(world)])
        assert (masks.sum(axis=0) == 1).all()
        for c in range(16):
            per_rank = masks[:, pc == c].sum(axis=1)
            assert per_rank.max() - per_rank.min() <= 1
