"""Timeline of the hand-overs inside the MF-SGD epoch kernel: python scripts/trace_sgd.py N cs max_run
(per-rank workload of an N-rank job on ONE GPU, see tune_sgd_virtual.py).  Prints, per step, how long the leaders
waited for their column-block and how long the cells took."""
import ctypes, sys
from pathlib import Path
import torch
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from harp_b200 import _native
from harp_b200.models.mfsgd import MFSGDEngine
N, cs, mr = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
dev = torch.device('cuda', 0)
rows, cols, nnz, rk = 480_000 // N, 17_700, 100_000_000 // N, 128
g = torch.Generator(device=dev).manual_seed(11)
r = (torch.rand(nnz, device=dev, generator=g) ** 2 * rows).long() % rows
c = (torch.rand(nnz, device=dev, generator=g) ** 2 * cols).long() % cols
v = torch.rand(nnz, device=dev, generator=g) * 4 + 1
kw = dict(virtual_world=N) if N > 1 else {}
eng = MFSGDEngine(r, c, v, rows, cols, rk, 0.001, 0.01, seed=3, cluster=cs, max_run=mr, fused=True, **kw)
for _ in range(3):
    eng.epoch()
torch.cuda.synchronize()
steps, B = max(N, 1) * eng.S, eng.B
tr = torch.zeros(steps * B * B * 27, dtype=torch.int64, device=dev)
lib = _native.lib()
lib.hb_sgd_epoch_trace.argtypes = [ctypes.c_void_p]
lib.hb_sgd_epoch_trace(ctypes.c_void_p(tr.data_ptr()))
eng.epoch(); torch.cuda.synchronize()
lib.hb_sgd_epoch_trace(ctypes.c_void_p(0))
full = tr.view(steps, B, B, 27).double().cpu()
t = full[..., :3]          # [step g][hand-over t][worker][wait0, go, done]
t0 = t[..., 0][t[..., 0] > 0].min()
t = (t - t0) / 1e3                                     # microseconds
wait = (t[..., 1] - t[..., 0]); cell = (t[..., 2] - t[..., 1])
print(f"N={N} cs={cs} B={B} max_run={eng.max_run}: epoch {float(t[..., 2].max()):.1f} us, {steps * B} hand-overs per worker")
print(f"per hand-over: wait mean {wait.mean():.2f} us (median {wait.median():.2f}, p95 {wait.flatten().quantile(0.95):.2f}), "
      f"cell mean {cell.mean():.2f} us (median {cell.median():.2f}, p95 {cell.flatten().quantile(0.95):.2f}, max {cell.max():.2f})")
print("per-worker busy share:", float(cell.sum() / (t[..., 2].max() * B)))
for gi in range(min(steps, 3)):
    print(f" step {gi}: start {float(t[gi, 0, :, 0].min()):.1f} end {float(t[gi, -1, :, 2].max()):.1f} cell max per hand-over:",
          [round(float(cell[gi, ti].max()), 1) for ti in range(min(B, 8))], "min:", [round(float(cell[gi, ti].min()), 1) for ti in range(min(B, 8))])

ncs = min(eng.cs, 8)                   # the trace slot records the first eight CTAs of a worker
c = (full[..., 3:3 + 3 * ncs].reshape(steps, B, B, ncs, 3) - t0) / 1e3      # per CTA: work start, work end, fenced
go = t[..., 1].unsqueeze(-1)
print(f"per CTA: go->work start {float((c[..., 0] - go).mean()):.2f} us, work {float((c[..., 1] - c[..., 0]).mean()):.2f} us "
      f"(max over CTAs of a worker {float((c[..., 1] - c[..., 0]).max(-1).values.mean()):.2f}), fence {float((c[..., 2] - c[..., 1]).mean()):.2f} us, "
      f"last CTA fenced -> leader done {float((t[..., 2] - c[..., 2].max(-1).values).mean()):.2f} us")
