"""Guard bands around every explicit CUDA allocation made while the context is active.

compute-sanitizer is closed on the GPU pool (profiles/sanitize.md), so out-of-bounds WRITES are caught this way:
`torch.empty / zeros / ones / full / empty_like / zeros_like` for CUDA tensors are routed to a byte buffer with a
512-byte sentinel band on each side; `check()` verifies every band that was ever handed out.  All output, scratch and
state buffers of the ops wrappers and engines are allocated with these functions, so a kernel that writes one element
past (or before) any of them trips the check."""
import contextlib
import math

import torch

BAND = 512
FILL = 0xA5


class _Guards:
    def __init__(self):
        self.bufs = []

    def alloc(self, shape, dtype, device, fill):
        n = int(math.prod(shape)) if len(shape) else 1
        esz = torch.empty((), dtype=dtype).element_size()
        nbytes = (n * esz + 255) // 256 * 256
        raw = self._orig_empty(nbytes + 2 * BAND, dtype=torch.uint8, device=device)
        raw[:BAND] = FILL
        raw[BAND + n * esz:] = FILL                      # the band starts right after the last element
        self.bufs.append((raw, n * esz))
        t = raw[BAND:BAND + n * esz].view(dtype).view(*shape) if n else self._orig_empty(*shape, dtype=dtype, device=device)
        if fill is not None and n:
            t.fill_(fill)
        return t

    def check(self):
        bad = []
        for i, (raw, nb) in enumerate(self.bufs):
            if not bool((raw[:BAND] == FILL).all()) or not bool((raw[BAND + nb:] == FILL).all()):
                bad.append((i, nb))
        return bad


def _shape_of(args, kwargs):
    if "size" in kwargs:
        return tuple(kwargs.pop("size"))
    if len(args) == 1 and isinstance(args[0], (tuple, list, torch.Size)):
        return tuple(int(x) for x in args[0])
    return tuple(int(x) for x in args)


@contextlib.contextmanager
def guard_bands(device_type: str = "cuda"):
    g = _Guards()
    orig = {n: getattr(torch, n) for n in ("empty", "zeros", "ones", "full", "empty_like", "zeros_like")}
    g._orig_empty = orig["empty"]

    def is_cuda(dev):
        return dev is not None and torch.device(dev).type == device_type

    def make(name, fill):
        def f(*args, **kwargs):
            dev = kwargs.get("device")
            if not is_cuda(dev) or kwargs.get("out") is not None or kwargs.get("pin_memory"):
                return orig[name](*args, **kwargs)
            kw = dict(kwargs)
            a = list(args)
            fv = fill
            if name == "full":
                fv = kw.pop("fill_value") if "fill_value" in kw else a.pop(1)
                shape = _shape_of(a[:1] if a else (), kw)
            else:
                shape = _shape_of(a, kw)
            dtype = kw.get("dtype")
            if dtype is None:
                dtype = (torch.bool if isinstance(fv, bool) else torch.int64 if isinstance(fv, int) else
                         torch.get_default_dtype()) if name == "full" else torch.get_default_dtype()
            return g.alloc(shape, dtype, torch.device(dev), fv)
        return f

    def make_like(name, fill):
        def f(t, **kwargs):
            dev = kwargs.get("device", t.device)
            if not is_cuda(dev) or kwargs.get("memory_format") not in (None, torch.contiguous_format, torch.preserve_format) or not t.is_contiguous():
                return orig[name](t, **kwargs)
            return g.alloc(tuple(t.shape), kwargs.get("dtype") or t.dtype, torch.device(dev), fill)
        return f

    torch.empty, torch.zeros, torch.ones = make("empty", None), make("zeros", 0), make("ones", 1)
    torch.full = make("full", None)
    torch.empty_like, torch.zeros_like = make_like("empty_like", None), make_like("zeros_like", 0)
    try:
        yield g
    finally:
        for n, fn in orig.items():
            setattr(torch, n, fn)
