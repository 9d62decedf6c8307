"""Every kernel family at CI-sized, tile-straddling shapes with sentinel bands around every buffer it is handed
(tests/guards.py): no launch may write outside what it was given.  Stands in for compute-sanitizer memcheck, which is
closed on the GPU pool (profiles/sanitize.md)."""
import importlib.util
import sys
from pathlib import Path

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = Path(__file__).resolve().parent.parent


def _targets():
    spec = importlib.util.spec_from_file_location("sanitize_targets", ROOT / "scripts" / "sanitize_targets.py")
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def test_the_guard_itself_catches_a_stray_write():
    if not torch.cuda.is_available():
        pytest.skip("needs a GPU")
    from guards import guard_bands
    with guard_bands() as g:
        t = torch.zeros(100, dtype=torch.float32, device="cuda")
        assert t.shape == (100,) and float(t.sum()) == 0.0
        u = torch.empty((7, 3), dtype=torch.int32, device="cuda")
        assert g.check() == []
        torch.as_strided(u, (22,), (1,))[21] = 5          # one element past the end of u
        assert [nb for _, nb in g.check()] == [84]


@pytest.mark.parametrize("section", ["kmeans", "gram", "mf", "lda", "linalg", "subgraph"])
def test_no_kernel_writes_outside_its_buffers(native_built, section):
    if not torch.cuda.is_available():
        pytest.skip("needs a GPU")
    sys.path.insert(0, str(Path(__file__).resolve().parent))
    from guards import guard_bands
    mod = _targets()
    with guard_bands() as g:
        mod.SECTIONS[section]()
        torch.cuda.synchronize()
        assert len(g.bufs) > 5, "the section allocated nothing through the guarded functions"
        assert g.check() == []
