"""One GPU, 1/8 of BASELINE config 3 (W is L2 resident): per-panel kernel vs epoch kernel at several worker sizes."""
import sys
from pathlib import Path
import torch
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from harp_b200.models.mfsgd import MFSGDEngine
dev = torch.device('cuda', 0)
rows, cols, nnz, rk = 60_000, 17_700, 12_500_000, 128
g = torch.Generator(device=dev).manual_seed(11)
r = (torch.rand(nnz, device=dev, generator=g) ** 2 * rows).long() % rows
c = (torch.rand(nnz, device=dev, generator=g) ** 2 * cols).long() % cols
v = torch.rand(nnz, device=dev, generator=g) * 4 + 1
for kw in (dict(fused=False), dict(fused=False, blocks=74), dict(fused=False, blocks=37), dict(fused=True, cluster=1), dict(fused=True, cluster=2), dict(fused=True, cluster=4), dict(fused=True, cluster=8)):
    eng = MFSGDEngine(r, c, v, rows, cols, rk, 0.001, 0.01, seed=3, **kw)
    for _ in range(3):
        eng.epoch()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        eng.epoch()
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 10
    print(kw, 'B', eng.B, 'cs', eng.cs, 'max_run', eng.max_run, f'ms/epoch {ms:.3f} -> {nnz / ms / 1e6:.2f} G/s', flush=True)
    del eng
