"""Developer aid: one emulated rank of a bucket-sharded index on the bench data, small n_buckets."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import torch
import bench
sys.argv = [sys.argv[0]] + sys.argv[1:]
rank = int(os.environ.get("EMU_RANK", "4")); world = int(os.environ.get("EMU_WORLD", "8"))
sys.argv = [sys.argv[0]]
args = bench.parse_args()
device = torch.device("cuda:0")
from lmi_b200.index import GpuLearnedIndex
inp = bench.build_inputs(args, device)
qn, qs = inp["queries_nav"].contiguous(), inp["queries_f16"].contiguous()
ranks = [int(x) for x in os.environ.get("EMU_RANKS", str(rank)).split(",")]
for rank in ranks:
    li = GpuLearnedIndex(args.categories, device=0, rank=rank, world=world)
    li.set_model(inp["weights"], inp["biases"])
    li.prepare_device(inp["data_f16"], inp["pred_categories"], id_base=1)
    points = [tuple(int(v) for v in pt.split(":")) for pt in os.environ.get("EMU_POINTS", "10:1,10:2,10:4,10:16,100:1,100:2,100:4,100:16").split(",")]
    reps = int(os.environ.get("EMU_REPS", "8"))
    for k, nb in points:
        if True:
            p = torch.empty(li.partial_bytes(args.nq, k), dtype=torch.uint8, device=device)
            for rep in range(reps):
                li.search_partial_device(qn, qs, nb, k, p)
            torch.cuda.synchronize()
            st = li.stats()
            print("ok", rank, k, nb, st["n_work_items"], st["chunk_objects"], st["pairs"], st["n_partial_lists"], flush=True)
    li.close()
