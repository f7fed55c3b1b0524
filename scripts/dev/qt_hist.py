"""Readers per object chunk of the bench workload: how many work items stream every chunk of the bucket store under the
CTA-pair kernel's decompositions, weighted by bytes -- the DRAM traffic K3 would have with no / perfect L2 sharing between
the equal-speed readers of a chunk.  python scripts/dev/qt_hist.py [--rows N] (one GPU)."""
import argparse, json, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows", type=int, default=10_000_000)
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "qt_hist.json"))
    a = ap.parse_args()
    import torch
    import bench
    sys.argv = [sys.argv[0], "--rows", str(a.rows)]
    args = bench.parse_args()
    device = torch.device("cuda", 0)
    inp = bench.build_inputs(args, device)
    from lmi_b200.index import GpuLearnedIndex
    li = GpuLearnedIndex(args.categories, device=0)
    li.set_model(inp["weights"], inp["biases"])
    C, nb = args.categories, args.buckets
    n_b = np.bincount(inp["pred_categories"].cpu().numpy(), minlength=C).astype(np.int64)
    _, cls = li.model.predict_proba(inp["queries_nav"].cpu().numpy(), n_top=nb)
    m_b = np.bincount(cls.reshape(-1), minlength=C).astype(np.int64)
    m_b[n_b == 0] = 0
    qt = (m_b + 127) // 128
    tot = float(n_b[m_b > 0].sum())
    hist = {int(t): float(n_b[qt == t].sum() / tot) for t in np.unique(qt)}
    full, odd = qt // 2, qt & 1
    readers_now = full + odd                                   # every pair item streams its chunk
    share_full = np.minimum(full, 1) + odd                     # full pairs of a chunk share one read, the half pair reads again
    share_full_q3 = np.where(qt == 3, 1, share_full)           # + three-tile buckets as three equal-speed half pairs
    share_all = (qt > 0).astype(np.int64)
    rep = dict(byte_share_by_query_tiles=hist,
               traffic_x_no_sharing=float((n_b * readers_now).sum() / tot),
               traffic_x_full_pairs_share=float((n_b * share_full).sum() / tot),
               traffic_x_full_pairs_share_and_q3_all_half=float((n_b * share_full_q3).sum() / tot),
               traffic_x_ideal=float((n_b * share_all).sum() / tot),
               tile_efficiency=float((m_b * n_b).sum() / (qt * 128 * n_b).sum()),
               mean_m_b=float(m_b[m_b > 0].mean()), buckets=int((m_b > 0).sum()))
    print(json.dumps(rep, indent=1))
    os.makedirs(os.path.dirname(a.out), exist_ok=True)
    json.dump(rep, open(a.out, "w"), indent=1)


if __name__ == "__main__":
    main()
