"""A/B of MF-SGD engine variants on the BASELINE config-3 shape (one GPU): python scripts/ab_sgd.py [fused-only]."""
import sys
from pathlib import Path
import torch
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from harp_b200.models.mfsgd import MFSGDEngine
dev = torch.device('cuda', 0)
rows, cols, nnz, rk = 480_000, 17_700, 100_000_000, 128
g = torch.Generator(device=dev).manual_seed(11)
r = (torch.rand(nnz, device=dev, generator=g) ** 2 * rows).long() % rows
c = (torch.rand(nnz, device=dev, generator=g) ** 2 * cols).long() % cols
v = torch.rand(nnz, device=dev, generator=g) * 4 + 1
variants = [dict(num_slices=2, fused=True), dict(num_slices=2, fused=False)]
if len(sys.argv) > 1 and sys.argv[1] == "all":
    variants += [dict(num_slices=1, fused=True), dict(num_slices=1, fused=False), dict(num_slices=2, fused=True, loopback=True)]
for kw in variants:
    eng = MFSGDEngine(r, c, v, rows, cols, rk, 0.001, 0.01, seed=3, **kw)
    for _ in range(2):
        eng.epoch()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(4):
        eng.epoch()
    e1.record(); torch.cuda.synchronize()
    print(kw, 'B', eng.B, 'cs', eng.cs, 'max_run', eng.max_run, 'ms/epoch', e0.elapsed_time(e1) / 4, flush=True)
    del eng
