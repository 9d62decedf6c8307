import sys, math, torch
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from harp_b200.models.mfsgd import MFSGDEngine
from harp_b200.ops import mf
dev = torch.device("cuda")
rows, cols, nnz, rk = 480_000, 17_700, 100_000_000, 128
g = torch.Generator(device=dev).manual_seed(11)
r = (torch.rand(nnz, device=dev, generator=g) ** 2 * rows).long() % rows
c = (torch.rand(nnz, device=dev, generator=g) ** 2 * cols).long() % cols
v = torch.rand(nnz, device=dev, generator=g) * 4 + 1
eng = MFSGDEngine(r, c, v, rows, cols, rk, 0.001, 0.01, seed=3)
print("B", eng.B, "S", eng.S, "max_blocks", mf.max_blocks(rk), [p.nnz for p in eng.panels[0]], [p.seg_col.numel() for p in eng.panels[0]])
def t(fn, n=3):
    for _ in range(2): fn()
    torch.cuda.synchronize()
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n
print("epoch ms", t(eng.epoch))
for k in range(2):
    p = eng.panels[0][k]
    print("slice", k, "pass ms", t(lambda: mf.sgd_pass(p, eng.W, eng.Hbuf[k][0], 0.001, 0.01)))
# same ratings as the kernel bench builds them (single slice with the same column count)
m = (c % 2) == 0
p2 = mf.build_panel(r[m], c[m] // 2, v[m], eng.B)
H2 = torch.rand(8850, rk, device=dev) / math.sqrt(rk)
print("rebuilt slice-0 panel pass ms", t(lambda: mf.sgd_pass(p2, eng.W, H2, 0.001, 0.01)))
W2 = torch.rand(rows, rk, device=dev) / math.sqrt(rk)
print("same, fresh W", t(lambda: mf.sgd_pass(p2, W2, H2, 0.001, 0.01)))
