"""Per-rank cost of a sharded search, measured on ONE GPU: an index built as rank r of `world` holds exactly the shard
that rank would hold on a real box, so navigation / plan / rescoring / finalize times of `lmi_search_partial_device` are
the real per-rank figures; only the exchange (all-gather + merge) needs several GPUs.  Development aid for tuning the
chunk rule and the fixed costs that bound strong scaling.

  python scripts/dev/shard_emulate.py --worlds 2,4,8 --chunks 0,8192,16384 [--policy buckets|stripes] [--k 10]
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--worlds", default="8")
    ap.add_argument("--ranks", default="0", help="comma list, or 'all'")
    ap.add_argument("--chunks", default="0")
    ap.add_argument("--policy", default="buckets")
    ap.add_argument("--k", type=int, default=10)
    ap.add_argument("--buckets", type=int, default=4)
    ap.add_argument("--rows", type=int, default=10_000_000)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--item-orders", default="0", help="comma list of item_order values (0 = full pairs first, 1 = bucket-major)")
    ap.add_argument("--locksteps", default="0", help="comma list of lockstep values (0 off, 1 on) tried for every order")
    ap.add_argument("--merge", action="store_true", help="also build every rank's shard, concatenate the partials (the all-gather) "
                                                         "and time merge_partials_kernel")
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "shard_emulate.json"))
    a = ap.parse_args()
    import torch
    import bench
    from lmi_b200.index import GpuLearnedIndex, SHARD_BUCKETS, SHARD_STRIPES
    sys.argv = [sys.argv[0], "--rows", str(a.rows), "--k", str(a.k), "--buckets", str(a.buckets)]
    args = bench.parse_args()
    dev = torch.device("cuda", 0)
    inp = bench.build_inputs(args, dev)
    qn, qs = inp["queries_nav"].contiguous(), inp["queries_f16"].contiguous()
    nq, k, nb = args.nq, a.k, a.buckets
    runs = []
    for world in [int(w) for w in a.worlds.split(",")]:
        ranks = range(world) if a.ranks == "all" else [int(r) for r in a.ranks.split(",") if int(r) < world]
        for rank in ranks:
            li = GpuLearnedIndex(args.categories, rank=rank, world=world,
                                 shard_policy=SHARD_BUCKETS if a.policy == "buckets" else SHARD_STRIPES)
            li.set_model(inp["weights"], inp["biases"])
            li.prepare_device(inp["data_f16"], inp["pred_categories"])
            part = torch.empty(li.partial_bytes(nq, k), dtype=torch.uint8, device=dev)
            ref_part = None
            for chunk, order, lock in [(int(c), int(o), int(l)) for c in a.chunks.split(",") for o in a.item_orders.split(",")
                                       for l in a.locksteps.split(",")]:
                li.set_option("chunk_objects", chunk)
                li.set_option("item_order", order)
                li.set_option("lockstep", lock)
                for _ in range(3):
                    li.search_partial_device(qn, qs, nb, k, part)
                torch.cuda.synchronize(); li.stats()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                for _ in range(a.steps):
                    li.search_partial_device(qn, qs, nb, k, part)
                e1.record(); torch.cuda.synchronize()
                st = li.stats()
                if ref_part is None:
                    ref_part = part.clone()
                same = bool(torch.equal(ref_part, part))
                runs.append(dict(world=world, rank=rank, policy=a.policy, k=k, item_order=order, lockstep=lock, partial_equal_to_first=same,
                                 chunk_objects=st["chunk_objects"],
                                 ms_partial_search=e0.elapsed_time(e1) / a.steps, ms_navigate=st["ms_navigate"], ms_plan=st["ms_plan"],
                                 ms_rescore_avg=st["ms_rescore_avg"], ms_finalize=st["ms_finalize"], kernel=st["rescore_kernel"],
                                 work_items=st["n_work_items"], partial_lists=st["n_partial_lists"], pairs=st["pairs"],
                                 tile_efficiency=st["pairs"] / max(1, st["padded_pairs"]),
                                 useful_tflops=2 * 768 * st["pairs"] / max(st["ms_rescore_avg"], 1e-6) / 1e9))
                print(json.dumps(runs[-1]), flush=True)
            li.close()
    if a.merge:
        for world in [int(w) for w in a.worlds.split(",")]:
            parts, keep = [], None
            fin = []
            for rank in range(world):
                li = GpuLearnedIndex(args.categories, rank=rank, world=world,
                                     shard_policy=SHARD_BUCKETS if a.policy == "buckets" else SHARD_STRIPES)
                li.set_model(inp["weights"], inp["biases"])
                li.prepare_device(inp["data_f16"], inp["pred_categories"])
                p_ = torch.empty(li.partial_bytes(nq, k), dtype=torch.uint8, device=dev)
                for _ in range(3):
                    li.search_partial_device(qn, qs, nb, k, p_)
                torch.cuda.synchronize()
                fin.append(li.stats()["ms_finalize"])
                parts.append(p_)
                if keep is None:
                    keep = li
                else:
                    li.close()
            gathered = torch.cat(parts)
            out = (torch.empty(nq, k, dtype=torch.float64, device=dev), torch.empty(nq, k, dtype=torch.int32, device=dev))
            for _ in range(3):
                keep.merge_partials_device(gathered, world, nq, k, out)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(a.steps):
                keep.merge_partials_device(gathered, world, nq, k, out)
            e1.record(); torch.cuda.synchronize()
            runs.append(dict(world=world, k=k, merge_ms=e0.elapsed_time(e1) / a.steps, finalize_ms_per_rank=fin,
                             gathered_mb=gathered.numel() / 1e6))
            print(json.dumps(runs[-1]), flush=True)
            keep.close()
    os.makedirs(os.path.dirname(a.out), exist_ok=True)
    json.dump(runs, open(a.out, "w"), indent=1)


if __name__ == "__main__":
    main()
