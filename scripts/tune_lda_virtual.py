"""Per-rank kernel time of LDA-CGS at N ranks, measured on ONE GPU: python scripts/tune_lda_virtual.py N [cs ...]
(LDAEngine(virtual_world=N): the rank's documents against N x 2 word-topic panels, loop-back ring)."""
import sys
from pathlib import Path
import torch
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from harp_b200.models.lda import LDAEngine
N = int(sys.argv[1]) if len(sys.argv) > 1 else 8
css = [int(x) for x in sys.argv[2:]] or [1, 2, 4, 8]
dev = torch.device('cuda', 0)
D, V, K, L = 1_000_000 // N, 100_000, 1000, 100
g = torch.Generator(device=dev).manual_seed(5)
docs = torch.arange(D, device=dev).repeat_interleave(L)
words = (torch.rand(D * L, device=dev, generator=g) ** 3 * V).long() % V
for cs in css:
    for mr in (None, 16):
        kw = dict(virtual_world=N) if N > 1 else {}
        eng = LDAEngine(docs, words, D, V, K, 0.01, 0.01, seed=1, cluster=cs, max_run=mr, **kw)
        for _ in range(2):
            eng.iterate()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(4):
            t = eng.iterate()
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 4
        ok = int(t.item()) == D * L
        print(f"N={N} cs={cs} B={eng.B} max_run={eng.max_run} fused={eng.fused} ms/iter {ms:.3f} -> {D * L / ms / 1e6:.3f} G tokens/s per GPU, "
              f"{N * D * L / ms / 1e6:.2f} G aggregate {'' if ok else 'TRAINED ' + str(int(t.item()))}", flush=True)
        del eng
