"""torchrun worker (gloo, CPU): the NumPy-facing HarpSession API (python/harp_session.py of the reference, working)."""
import os
import sys
from pathlib import Path

import numpy as np

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
os.environ["HARP_FORCE_CPU"] = "1"
from harp_b200.partition import Operation
from harp_b200.session import HarpSession, PartitioningMode, Type


def main():
    s = HarpSession("demo")
    me, n = s.ctx.self_id, s.ctx.num_workers
    assert s.name == "demo" and s.com.barrier("demo", "b0")
    a = np.full(5, float(me + 1))
    assert s.com.broadcast("demo", "bc", a, Type.DOUBLE, PartitioningMode.HOMOGENEOUS, 1 % n, False)
    np.testing.assert_array_equal(a, np.full(5, float(1 % n + 1)))
    b = np.arange(6, dtype=np.float32).reshape(2, 3) * (me + 1)
    assert s.com.allreduce("demo", "ar", b, Type.FLOAT)
    np.testing.assert_allclose(b, np.arange(6, dtype=np.float32).reshape(2, 3) * (n * (n + 1) / 2))
    c = np.array([me, 10 - me], dtype=np.int64)
    assert s.com.allreduce("demo", "mx", c, Type.LONG, op=Operation.MAX)
    np.testing.assert_array_equal(c, np.array([n - 1, 10]))
    d = np.full(3, float(me + 1))
    assert s.com.reduce("demo", "rd", d, Type.DOUBLE, 0)
    if me == 0:
        np.testing.assert_array_equal(d, np.full(3, n * (n + 1) / 2))
    else:
        np.testing.assert_array_equal(d, np.full(3, float(me + 1)))
    g = s.com.allgather("demo", "ag", np.array([[me, me * 2]], dtype=np.int32), Type.INT)
    assert g.shape == (n, 1, 2) and [int(x) for x in g[:, 0, 1]] == [2 * r for r in range(n)]
    try:
        s.com.allreduce("demo", "bad", np.zeros(2, dtype=np.float32), Type.DOUBLE)
        raise SystemExit("dtype mismatch not detected")
    except TypeError:
        pass
    s.com.barrier("demo", "b1")
    if me == 0:
        print(f"SESSION_OK {n}", flush=True)


if __name__ == "__main__":
    main()
