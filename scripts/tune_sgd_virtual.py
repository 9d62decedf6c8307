"""Per-rank kernel time of MF-SGD at N ranks, measured on ONE GPU: python scripts/tune_sgd_virtual.py N [cs ...].
The rank's share of BASELINE config 3 (rows / N, all columns, ratings / N) is laid out as N x 2 panels and run through
the whole-epoch kernel with a loop-back ring (MFSGDEngine(virtual_world=N))."""
import sys
from pathlib import Path
import torch
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from harp_b200.models.mfsgd import MFSGDEngine
N = int(sys.argv[1]) if len(sys.argv) > 1 else 8
css = [int(x) for x in sys.argv[2:]] or [1, 2, 4, 8]
dev = torch.device('cuda', 0)
rows, cols, nnz, rk = 480_000 // N, 17_700, 100_000_000 // N, 128
g = torch.Generator(device=dev).manual_seed(11)
r = (torch.rand(nnz, device=dev, generator=g) ** 2 * rows).long() % rows
c = (torch.rand(nnz, device=dev, generator=g) ** 2 * cols).long() % cols
v = torch.rand(nnz, device=dev, generator=g) * 4 + 1
for cs in css:
    import os
    for extra in [int(x) for x in os.environ.get("TUNE_EXTRA_C", "0").split(",")]:
        mr = int(os.environ["TUNE_MAX_RUN"]) if os.environ.get("TUNE_MAX_RUN") else None
        eng = MFSGDEngine(r, c, v, rows, cols, rk, 0.001, 0.01, seed=3, virtual_world=N, cluster=cs, max_run=mr,
                          col_blocks=(None if extra == 0 else -extra))
        for _ in range(3):
            eng.epoch()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(10):
            t = eng.epoch()
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 10
        words = eng.ring.state[:64].view(torch.int32).cpu().tolist()
        if int(t.item()) != nnz:
            print('TRAINED', int(t.item()), 'of', nnz, 'ctl', words[:12], flush=True)
        print(f"N={N} cs={cs} B={eng.B} C={eng.C} max_run={eng.max_run} ms/epoch {ms:.3f} -> {nnz / ms / 1e6:.2f} G updates/s per GPU, {N * nnz / ms / 1e6:.1f} G aggregate", flush=True)
        del eng
