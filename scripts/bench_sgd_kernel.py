"""Kernel-level throughput of the SGD pass (Netflix-shaped synthetic panel)."""
import json, sys, math
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
from harp_b200.ops import mf

nnz = int(sys.argv[1]) if len(sys.argv) > 1 else 20_000_000
nr = int(sys.argv[2]) if len(sys.argv) > 2 else 480_000
nc = int(sys.argv[3]) if len(sys.argv) > 3 else 17_700
rank = int(sys.argv[4]) if len(sys.argv) > 4 else 128
g = torch.Generator(device="cuda").manual_seed(0)
r = (torch.rand(nnz, device="cuda", generator=g) ** 2 * nr).long() % nr
c = (torch.rand(nnz, device="cuda", generator=g) ** 2 * nc).long() % nc
v = torch.rand(nnz, device="cuda", generator=g) * 4 + 1
cmuls = [float(x) for x in (sys.argv[6].split(",") if len(sys.argv) > 6 else ["1"])]
import itertools
for B, cm in itertools.product([int(x) for x in (sys.argv[5].split(",") if len(sys.argv) > 5 else ["148", "296", "444"])], cmuls):
    B = min(B, mf.max_blocks(rank))
    panel = mf.build_panel(r, c, v, B, C=int(B * cm))
    W = torch.rand(nr, rank, device="cuda") / math.sqrt(rank); H = torch.rand(nc, rank, device="cuda") / math.sqrt(rank)
    for _ in range(2): mf.sgd_pass(panel, W, H, 0.001, 0.01)
    torch.cuda.synchronize()
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(3): mf.sgd_pass(panel, W, H, 0.001, 0.01)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 3
    seg = panel.seg_col.numel()
    print(json.dumps({"B": B, "C": panel.ncolblocks, "nnz": nnz, "rank": rank, "ms_per_pass": ms, "updates_per_s": nnz / ms * 1e3,
                      "avg_run_len": nnz / seg, "GBps_min": nnz * (2 * rank * 4) / ms / 1e6}))
