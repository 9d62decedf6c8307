"""Times one colouring of the colour-coding DP on one GPU: python scripts/bench_subgraph.py [n] [deg] [template size]
(random graph, path-with-branches template).  Prints one JSON line: ms per colouring, edge-visits/s, and the time of the
same DP through torch.sparse.mm + gathers on 2^k-wide fp64 tables (the round-1 formulation) on a smaller slice."""
import json
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch

from harp_b200.apps.graph import SubTemplate, colour_coding_pass

n = int(sys.argv[1]) if len(sys.argv) > 1 else 2_000_000
deg = int(sys.argv[2]) if len(sys.argv) > 2 else 8
k = int(sys.argv[3]) if len(sys.argv) > 3 else 10
dev = torch.device("cuda:0")
g = torch.Generator(device=dev).manual_seed(1)
src = torch.arange(n, device=dev).repeat_interleave(deg)
dst = torch.randint(0, n, (n * deg,), device=dev, generator=g)
a = torch.cat([src, dst]); b = torch.cat([dst, src])
key = torch.unique(a * n + b)
a, b = key // n, key % n
keep = a != b
a, b = a[keep], b[keep]
ptr = torch.zeros(n + 1, dtype=torch.int64, device=dev)
ptr[1:] = torch.cumsum(torch.bincount(a, minlength=n), 0)
idx = b.to(torch.int32)
edges = [(i, i + 1) for i in range(k // 2)] + [(0, k // 2 + 1)] + [(i, i + 1) for i in range(k // 2 + 1, k - 1)]
adj_t = {i: [] for i in range(k)}
for x, y in edges:
    adj_t[x].append(y); adj_t[y].append(x)
tmpl = SubTemplate(0, frozenset(range(k)), adj_t)
colours = torch.randint(0, k, (n,), device=dev, generator=g)
for _ in range(2):
    colour_coding_pass(ptr, idx, colours, tmpl, k)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
steps = 5
for _ in range(steps):
    tab = colour_coding_pass(ptr, idx, colours, tmpl, k)
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / steps
print(json.dumps({"kernel": "subgraph_colour_coding", "vertices": n, "edges_directed": int(idx.numel()), "template_size": k,
                  "ms_per_colouring": round(ms, 3), "edge_visits_per_s": round(int(idx.numel()) * (k - 1) / ms * 1e3),
                  "colourful_count": float(tab[:, 0].double().sum())}))
