"""A/B of the two TMEM-resident rescoring kernels on one index: rescore_ts_kernel (option rescore = 1, default)
vs rescore_ts2_kernel (rescore = 3, tcgen05 cta_group::2 CTA pairs).  Development aid, not a bench line.

  stage 1 (--diag): one bucket, 256 queries -> one CTA pair; mismatches are broken down by query half
                    (leader / peer CTA) and by object position inside the 64-object tile (leader's / peer's
                    half of the B operand), which tells apart an operand-orientation error from a protocol bug.
  stage 2          : bench.py's 10M workload built once; both kernels alternately, K3 time / clock / padded
                    MMA work from lmi_get_stats, results compared bit for bit.

Usage: python scripts/dev/pair_ab.py [--rows 10000000] [--steps 10] [--out gpurun_out/pair_ab.json]
"""
import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)


def diag(nq=256, sets=0):
    import torch
    from lmi_b200.index import GpuLearnedIndex
    rng = np.random.default_rng(7)
    n, d = 4096 + 40, 768
    x = rng.standard_normal((n, d)).astype(np.float32)
    x /= np.linalg.norm(x, axis=1, keepdims=True)
    qv = x[rng.integers(0, n, nq)] + 0.3 * rng.standard_normal((nq, d)).astype(np.float32)
    x16, q16 = x.astype(np.float16), qv.astype(np.float16)
    pc = np.zeros(n, dtype=np.int64)
    cat = np.zeros(nq, dtype=np.int64)
    out = {}
    res = {}
    for name, r in (("ts", 1), ("ts2", 3)):
        li = GpuLearnedIndex(1, d=d, rescore=r, chunk_objects=2048)
        li.set_option("epi_sets", sets)
        li.prepare(x16, pc)
        dd, aa = li.search_single(None, x16, q16, cat, k=10)
        st = li.stats()
        res[name] = (dd, aa)
        out[name] = dict(kernel=st["rescore_kernel"], items=st["n_work_items"], uncertified=st["n_uncertified"])
        li.close()
    (d1, a1), (d3, a3) = res["ts"], res["ts2"]
    bad = (a1 != a3).any(axis=1)
    out["rows_mismatch"] = int(bad.sum())
    out["rows_mismatch_leader_half"] = int(bad[:128].sum())
    out["rows_mismatch_peer_half"] = int(bad[128:].sum())
    # ids are 1-based row numbers; one bucket -> CSR position = id - 1; chunks of 2048 start on multiples of 64
    missing = [int(i) for r in np.where(bad)[0] for i in a1[r] if i not in set(a3[r].tolist())]
    pos = (np.array(missing, dtype=np.int64) - 1) % 64
    out["missing_ids_in_leader_half_of_tile"] = int((pos < 32).sum())
    out["missing_ids_in_peer_half_of_tile"] = int((pos >= 32).sum())
    out["example_default"] = a1[0].tolist()
    out["example_pair"] = a3[0].tolist()
    out["dists_equal"] = bool((d1 == d3).all())
    torch.cuda.synchronize()
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows", type=int, default=10_000_000)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "pair_ab.json"))
    ap.add_argument("--diag-only", action="store_true")
    ap.add_argument("--chunks", default="0,16128,64000", help="chunk_objects values tried for the pair kernel (0 = automatic)")
    ap.add_argument("--dbg", default="", help="comma list of dbg_flags run for BOTH kernels after the A/B (results invalid by construction)")
    ap.add_argument("--skip-diag", action="store_true")
    ap.add_argument("--seq", default="", help="explicit run sequence 'rescore:chunk:flags,...' (replaces the default A/B order; "
                                              "interleave variants -- the power-capped clock drifts over a process lifetime)")
    ap.add_argument("--dbg-pair-only", action="store_true", help="run the --dbg flags on the pair kernel only")
    ap.add_argument("--diag-sets", type=int, default=0, help="epi_sets used by the diagnosis stage (0 = automatic)")
    ap.add_argument("--buckets", type=int, default=0, help="n_buckets of the searches (0 = bench default)")
    ap.add_argument("--exp", action="store_true", help="load liblmi_b200_exp.so (-DLMI_EXPERIMENTS build: dbg_flags honoured)")
    a = ap.parse_args()
    if a.exp:
        import lmi_b200.build as b
        from lmi_b200 import _lib
        _lib.LIB_PATH = b.build_experiments()
    os.makedirs(os.path.dirname(a.out), exist_ok=True)
    report = {}
    if not a.skip_diag:
        report["diag"] = diag(256, a.diag_sets)
        print(json.dumps(report["diag"]), flush=True)
        json.dump(report, open(a.out, "w"), indent=1)
        report["diag_odd_tile"] = diag(300, a.diag_sets)          # 3 query tiles: a full pair + a split odd tile
        print(json.dumps(report["diag_odd_tile"]), flush=True)
        json.dump(report, open(a.out, "w"), indent=1)
        if a.diag_only or report["diag"]["rows_mismatch"] or report["diag_odd_tile"]["rows_mismatch"]:
            return
    import torch
    import bench
    sys.argv = [sys.argv[0], "--rows", str(a.rows)] + (["--buckets", str(a.buckets)] if a.buckets else [])
    args = bench.parse_args()
    device = torch.device("cuda", 0)
    inp = bench.build_inputs(args, device)
    from lmi_b200.index import GpuLearnedIndex
    li = GpuLearnedIndex(args.categories, device=0, rescore=1)
    li.set_model(inp["weights"], inp["biases"])
    li.prepare_device(inp["data_f16"], inp["pred_categories"], id_base=1)
    qn, qs = inp["queries_nav"].contiguous(), inp["queries_f16"].contiguous()
    nq, k, nb = args.nq, args.k, args.buckets
    runs = []
    ref = None
    configs = [(1, 0, 0)] + [(3, int(c), 0) for c in a.chunks.split(",") if c != ""] + [(1, 0, 0)]
    for f in [int(v) for v in a.dbg.split(",") if v != ""]:
        configs += [(3, 0, f)] if a.dbg_pair_only else [(1, 0, f), (3, 0, f)]
    if a.seq:
        configs = [tuple(int(v) for v in item.split(":")) for item in a.seq.split(",")]       # rescore:chunk:flags[:epi_sets[:item_order]]
    for cfg in configs:
        rescore, chunk, flags = cfg[:3]
        sets = cfg[3] if len(cfg) > 3 else 0
        order = cfg[4] if len(cfg) > 4 else 0
        hint = cfg[5] if len(cfg) > 5 else 0
        lock = cfg[6] if len(cfg) > 6 else 0
        li.set_option("lockstep", lock)
        li.set_option("item_order", order)
        li.set_option("l2_hint", hint)
        li.set_option("rescore", rescore)
        li.set_option("chunk_objects", chunk)
        li.set_option("epi_sets", sets)
        if a.exp or flags:
            li.set_option("dbg_flags", flags)
        out = (torch.empty(nq, k, dtype=torch.float64, device=device), torch.empty(nq, k, dtype=torch.int32, device=device))
        for _ in range(3):
            li.search_device(qn, qs, nb, k, out)
        torch.cuda.synchronize()
        li.stats()
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ev0.record()
        for _ in range(a.steps):
            li.search_device(qn, qs, nb, k, out)
        ev1.record()
        torch.cuda.synchronize()
        st = li.stats()
        ids = out[1].cpu().numpy()
        dd = out[0].cpu().numpy()
        if ref is None:
            ref = (dd, ids)
        runs.append(dict(rescore=rescore, dbg_flags=flags, epi_sets=sets, item_order=order, l2_hint=hint, lockstep=lock, kernel=st["rescore_kernel"], chunk_objects=st["chunk_objects"],
                         ms_per_search=ev0.elapsed_time(ev1) / a.steps, k3_ms_avg=st["ms_rescore_avg"],
                         k3_sm_mhz=st["k3_sm_mhz"], work_items=st["n_work_items"], pairs=st["pairs"],
                         padded_pairs=st["padded_pairs"], tile_efficiency=st["pairs"] / max(st["padded_pairs"], 1),
                         useful_tflops=2 * 768 * st["pairs"] / (st["ms_rescore_avg"] * 1e-3) / 1e12,
                         ms_finalize=st["ms_finalize"], n_uncertified=st["n_uncertified"],
                         ids_equal_to_first=bool((ids == ref[1]).all()), dists_equal_to_first=bool((dd == ref[0]).all())))
        print(json.dumps(runs[-1]), flush=True)
        report["runs"] = runs
        json.dump(report, open(a.out, "w"), indent=1)


if __name__ == "__main__":
    main()
