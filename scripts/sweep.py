#!/usr/bin/env python
"""Recall-vs-QPS sweep of the query path (BASELINE.json configs[4], SURVEY.md 8(d) "10M sweep").

    python scripts/sweep.py [--n 10000000] [--buckets 1,2,3,4,6,8,12,16] [--ks 10,100] [--out gpurun_out/sweep]

Same synthetic inputs, model and index as ``bench.py`` (its ``build_inputs`` is reused); for every
(k, n_buckets) point: QPS with inputs resident in HBM (CUDA events, 3 warm-ups, --steps timed
searches), recall@k against the independent brute force, the rescoring kernel's average duration
and the roofline that bounds it (tensor or HBM).  The reference calls
``li.search(n_buckets=...)`` once per stop-condition value (search/search.py:115-128); so does this.
Writes ``<out>.jsonl`` (one line per point) and ``<out>.md`` (table).  Needs a B200.
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import bench  # noqa: E402


def main():
    p = argparse.ArgumentParser()
    p.add_argument("--n", type=int, default=10_000_000)
    p.add_argument("--nq", type=int, default=10_000)
    p.add_argument("--buckets", default="1,2,3,4,6,8,12,16")
    p.add_argument("--ks", default="10,100")
    p.add_argument("--steps", type=int, default=5)
    p.add_argument("--out", default="gpurun_out/sweep")
    a = p.parse_args()
    sys.argv = [sys.argv[0], "--n", str(a.n), "--nq", str(a.nq)]
    args = bench.parse_args()

    import torch
    if not torch.cuda.is_available():
        raise SystemExit("sweep.py: no CUDA device; the lmi_b200 path has no CPU fallback")
    world, rank = int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    device = torch.device("cuda", local_rank)
    torch.cuda.set_device(device)
    dist = None
    if world > 1:                                  # torchrun: bucket-sharded, one all-gather per search
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=device)
    from lmi_b200.index import GpuLearnedIndex
    from lmi_b200.shard import DistributedSearch
    peaks = bench.load_peaks()
    inp = bench.build_inputs(args, device, dist)
    li = GpuLearnedIndex(args.categories, device=local_rank, rank=rank, world=world)
    li.set_model(inp["weights"], inp["biases"])
    li.prepare_device(inp["data_f16"], inp["pred_categories"], id_base=1)
    ds = DistributedSearch(li)
    qn, qs = inp["queries_nav"].contiguous(), inp["queries_f16"].contiguous()
    nq = a.nq
    ks = [int(x) for x in a.ks.split(",")]
    nbs = [int(x) for x in a.buckets.split(",")]
    truth = bench.ground_truth(inp["data_f16"], qs, max(ks))          # one brute force, prefix per k
    if rank != 0:
        import tempfile
        a.out = os.path.join(tempfile.gettempdir(), f"sweep_rank{rank}")     # only rank 0's files are kept

    os.makedirs(os.path.dirname(a.out) or ".", exist_ok=True)
    rows = []
    with open(a.out + ".jsonl", "w") as fh:
        for k in ks:
            out = (torch.empty(nq, k, dtype=torch.float64, device=device), torch.empty(nq, k, dtype=torch.int32, device=device))
            for nb in nbs:
                for _ in range(3):
                    ds.search_device(qn, qs, nb, k, out)
                if dist:
                    dist.barrier()
                torch.cuda.synchronize()
                li.stats()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                for _ in range(a.steps):
                    ds.search_device(qn, qs, nb, k, out)
                e1.record()
                torch.cuda.synchronize()
                ms = e0.elapsed_time(e1) / a.steps
                if dist:                               # max over ranks
                    tm = torch.tensor([ms], device=device)
                    dist.all_reduce(tm, op=dist.ReduceOp.MAX)
                    ms = float(tm.item())
                st = li.stats()
                flops = 2.0 * li.d * st["pairs"]
                bytes_algo = st["bucket_bytes"] * (1 + 2.0 / li.d) + 2.0 * li.d * nq + 8.0 * nq * st["kprime"]
                t_gemm, t_hbm = flops / (peaks["tflops"] * 1e12), bytes_algo / (peaks["hbm_gbs"] * 1e9)
                t_k = st["ms_rescore_avg"] * 1e-3
                rec = bench.recall_at_k(out[1], truth, k)
                row = dict(n_gpus=world, k=k, n_buckets=nb, qps=nq / (ms * 1e-3), ms_per_search=ms, recall=rec,
                           rescore_ms=t_k * 1e3, kernel=st["rescore_kernel"], bound="tensor" if t_gemm >= t_hbm else "hbm",
                           t_roof_ms=max(t_gemm, t_hbm) * 1e3, roofline_frac=max(t_gemm, t_hbm) / t_k,
                           tflops=flops / t_k / 1e12, hbm_gbs=bytes_algo / t_k / 1e9,
                           n_uncertified=st["n_uncertified"], work_items=st["n_work_items"], chunk_objects=st["chunk_objects"])
                rows.append(row)
                fh.write(json.dumps(row) + "\n"); fh.flush()
                if rank == 0:
                    print(json.dumps(row), flush=True)
    with open(a.out + ".md", "w") as fh:
        fh.write(f"Sweep on {a.n} x 768, C={args.categories}, nq={nq}, sigma={args.sigma}, {world} GPU(s) (K3 columns: rank 0's shard; burst peaks: "
                 f"{peaks['tflops']:.0f} TFLOP/s, {peaks['hbm_gbs']:.0f} GB/s)\n\n")
        fh.write("| k | n_buckets | recall@k | QPS | search ms | K3 ms | bound | T_roof ms | frac of roofline | TFLOP/s | GB/s | uncertified |\n")
        fh.write("|---|---|---|---|---|---|---|---|---|---|---|---|\n")
        for r in rows:
            fh.write(f"| {r['k']} | {r['n_buckets']} | {r['recall']:.4f} | {r['qps']:,.0f} | {r['ms_per_search']:.2f} | "
                     f"{r['rescore_ms']:.2f} | {r['bound']} | {r['t_roof_ms']:.2f} | {r['roofline_frac']:.2f} | "
                     f"{r['tflops']:.0f} | {r['hbm_gbs']:.0f} | {r['n_uncertified']} |\n")


    if dist:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
