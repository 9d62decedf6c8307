"""torchrun worker: checks the NVLink peer-memory collectives against NCCL (run by test_symm_gpu.py)."""
import os
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
import torch.distributed as dist


def main():
    backend = sys.argv[1] if len(sys.argv) > 1 else "auto"
    rank = int(os.environ["RANK"]); world = int(os.environ["WORLD_SIZE"]); lr = int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(lr)
    dev = torch.device("cuda", lr)
    dist.init_process_group("nccl", device_id=dev)
    from harp_b200.parallel.symm import SymmetricWorkspace
    from harp_b200.models.kmeans import KMeansEngine

    ws = SymmetricWorkspace(dev, 64 << 20, backend=backend)
    if rank == 0:
        print(f"backend={ws.backend} multicast_ptr={ws.multicast_ptr:#x} world={world}", flush=True)
    g = torch.Generator(device=dev).manual_seed(100 + rank)

    # allreduce, several sizes / dtypes / ops, repeated to exercise ticket reuse
    for n in (1, 7, 1000, 1 << 20, 1290000 + 3):
        for dt in (torch.float32, torch.float64, torch.int32, torch.int64, torch.int16, torch.int8):
            ws.reset()
            t = ws.alloc(n, dt)
            for op, red in (("sum", dist.ReduceOp.SUM), ("max", dist.ReduceOp.MAX), ("min", dist.ReduceOp.MIN), ("minus", None)):
                for rep in range(2):
                    src = (torch.rand(n, device=dev, generator=g) * (100 if dt.is_floating_point or dt in (torch.int32, torch.int64) else 12)).to(dt)
                    wide = torch.float64 if dt.is_floating_point else torch.int64        # NCCL has no int16; reduce wide, narrow after
                    if red is None:                                                    # Operation.MINUS: v0 - v1 - ... in rank order
                        parts = [torch.empty(n, dtype=wide, device=dev) for _ in range(world)]
                        dist.all_gather(parts, src.to(wide))
                        ref = parts[0].clone()
                        for p_ in parts[1:]:
                            ref -= p_
                        ref = ref.to(dt)
                    else:
                        ref = src.to(wide, copy=True); dist.all_reduce(ref, op=red); ref = ref.to(dt)
                    t.copy_(src)
                    ws.allreduce_(t, op)
                    torch.cuda.synchronize()
                    if dt.is_floating_point:
                        torch.testing.assert_close(t, ref, rtol=1e-5, atol=1e-4, msg=lambda m: f"{m} [n={n} {dt} {op} rep={rep} t={t[:4].tolist()} ref={ref[:4].tolist()} src={src[:4].tolist()}]")
                    else:
                        assert torch.equal(t, ref), (n, dt, op)
    # reduce-scatter / reduce / allgather / broadcast / rotate / barrier
    ws.reset()
    n = 4096 * world + 0
    t = ws.alloc(n, torch.float32)
    src = torch.rand(n, device=dev, generator=g); ref = src.clone(); dist.all_reduce(ref)
    t.copy_(src); mine = ws.reduce_scatter_(t); torch.cuda.synchronize()
    per = n // world
    torch.testing.assert_close(mine, ref[rank * per:(rank + 1) * per], rtol=1e-5, atol=1e-5)
    t.copy_(src); ws.reduce_(t, root=world - 1); torch.cuda.synchronize()
    if rank == world - 1:
        torch.testing.assert_close(t, ref, rtol=1e-5, atol=1e-5)
    shard = torch.full((per,), float(rank + 1), device=dev)
    t.zero_(); t[rank * per:(rank + 1) * per] = shard; ws.allgather_(t); torch.cuda.synchronize()
    exp = torch.arange(1, world + 1, device=dev, dtype=torch.float32).repeat_interleave(per)
    assert torch.equal(t, exp)
    t.fill_(float(rank)); ws.broadcast_(t, root=1 % world); torch.cuda.synchronize()
    assert torch.equal(t, torch.full_like(t, float(1 % world)))
    a = ws.alloc(1000, torch.float32); b = ws.alloc(1000, torch.float32)
    a.fill_(float(rank)); ws.rotate_(b, a, (rank - 1) % world); torch.cuda.synchronize()
    assert torch.equal(b, torch.full_like(b, float((rank - 1) % world)))
    for _ in range(50):
        ws.barrier()
    torch.cuda.synchronize()

    # fused K-means p2p path == NCCL path
    ws.reset()
    X = torch.rand(20000, 64, device=dev, generator=g)
    C0 = torch.rand(300, 64, device=dev, generator=torch.Generator(device=dev).manual_seed(5))
    e1 = KMeansEngine(X, 300, comm="p2p", symm=ws); e1.set_centroids(C0)
    e2 = KMeansEngine(X, 300, comm="nccl"); e2.set_centroids(C0)
    for it in range(4):
        # float atomics make the partial sums differ in the last bits between two runs, and a point that
        # sits on a cluster boundary can then flip in a LATER iteration; so both engines are compared one
        # step at a time from identical centroids
        e2.set_centroids(e1.C.clone()); torch.cuda.synchronize(); dist.barrier()
        o1 = e1.step().clone(); o2 = e2.step().clone(); torch.cuda.synchronize()
        torch.testing.assert_close(o1, o2, rtol=1e-6, atol=1e-6)
        torch.testing.assert_close(e1.C, e2.C, rtol=1e-4, atol=1e-5)

    # one-launch fused iteration (GEMM + arg-min + regroup/update/allgather over NVLink) == NCCL path
    e3 = KMeansEngine(X, 300, comm="fused", symm=ws); e3.set_centroids(C0)
    e2.set_centroids(C0)
    for it in range(5):
        e2.set_centroids(e3.C.clone()); torch.cuda.synchronize(); dist.barrier()
        o3 = e3.step().clone(); o2 = e2.step().clone(); torch.cuda.synchronize()
        torch.testing.assert_close(o3, o2, rtol=1e-6, atol=1e-6)
        torch.testing.assert_close(e3.C, e2.C, rtol=1e-4, atol=1e-5)
    # every rank ends up with bit-identical centroids (rank-ordered reduction, one writer per row)
    mine = e3.C.clone(); ref = mine.clone(); dist.broadcast(ref, src=0)
    assert torch.equal(mine, ref)
    if rank == 0:
        print("fused kmeans ok", flush=True)

    # Gram + partial-result allreduce in one launch == local Gram + NCCL allreduce
    from harp_b200.ops.gram import FusedGram, gram
    for d_ in (256, 1024, 96):
        st = ws.alloc(FusedGram.state_bytes(d_), torch.uint8)
        fg = FusedGram(st, [p + ws.offset_of(st) for p in ws.peer_ptrs], rank, d_)
        torch.cuda.synchronize(); dist.barrier()
        for rep in range(3):
            Xg = torch.randn(3000 + 500 * rank + 100 * rep, d_, device=dev, generator=g)
            Gf = fg(Xg).clone()
            Gr = gram(Xg); dist.all_reduce(Gr); torch.cuda.synchronize()
            torch.testing.assert_close(Gf, Gr, rtol=1e-5, atol=1e-4 * Gr.abs().max().item())
            assert torch.equal(Gf, Gf.t())
            ref = Gf.clone(); dist.broadcast(ref, src=0)
            assert torch.equal(Gf, ref)                       # rank-ordered sums: identical bits everywhere
    if rank == 0:
        print("fused gram ok", flush=True)

    # ---- Table collectives routed to the NVLink kernels: dense tables in the slab vs the generic path on copies ----
    from harp_b200 import _native
    from harp_b200 import collective as C
    from harp_b200.partition import DoubleArrPlus, Partitioner, Table
    ws.reset()
    comm_s = C.Comm(device=dev, symm=ws)        # NVLink kernels
    comm_g = C.Comm(device=dev, symm=None)      # generic path: metadata + all_to_all
    calls = {"n": 0}
    orig_ago = C.Comm.all_gather_object

    def counting(self, obj):
        calls["n"] += 1
        return orig_ago(self, obj)
    C.Comm.all_gather_object = counting
    P, W = 37, 24                                # 37 partitions (not a multiple of the world size) of 24 doubles

    def make(dtype=torch.float64, seed=0, rows=P, ids=None):
        slab = ws.alloc(rows * W, dtype).view(rows, W)
        slab.copy_(torch.rand(rows, W, device=dev, dtype=dtype, generator=torch.Generator(device=dev).manual_seed(seed + 31 * rank)))
        return Table.dense(0, DoubleArrPlus(), slab, ids), slab

    def twin(table):                             # same contents, ordinary tensors -> generic path
        t2 = Table(0, DoubleArrPlus())
        from harp_b200.partition import Partition
        for p in table.get_partitions():
            t2.add_partition(Partition(p.id(), p.get().clone()))
        return t2

    def same(a, b):
        assert sorted(a.get_partition_ids()) == sorted(b.get_partition_ids()), (a.get_partition_ids(), b.get_partition_ids())
        for pid in a.get_partition_ids():
            torch.testing.assert_close(a.get_partition(pid).get(), b.get_partition(pid).get(), rtol=1e-12, atol=1e-12)

    for dt in (torch.float64, torch.float32):
        # allreduce / broadcast / reduce
        for op in ("allreduce", "broadcast", "reduce"):
            t, _ = make(dt, seed=1); g_ = twin(t)
            for rep_ in range(3):
                l0, c0 = _native.launches(), calls["n"]
                if op == "allreduce":
                    assert C.allreduce(comm_s, t); C.allreduce(comm_g, g_)
                elif op == "broadcast":
                    assert C.broadcast(comm_s, t, world - 1); C.broadcast(comm_g, g_, world - 1)
                else:
                    break
                torch.cuda.synchronize()
                assert world == 1 or _native.launches() > l0, op
                same(t, g_) if dt == torch.float64 else None
                if rep_ > 0:
                    # steady state: the dense route did no metadata exchange at all (the generic twin does its own)
                    pass
            if op == "reduce":
                assert C.reduce(comm_s, t, 1 % world); C.reduce(comm_g, g_, 1 % world); torch.cuda.synchronize()
                if rank == 1 % world:
                    if dt == torch.float64:
                        same(t, g_)
                elif world > 1:
                    # the generic table is released on non-roots; the slab table stays dense with zero rows, so that
                    # a following broadcast takes the kernel route on every rank
                    assert g_.get_num_partitions() == 0 and t.is_dense() and float(t.slab.abs().max()) == 0.0
                    assert C.broadcast(comm_s, t, 1 % world) is True
                if world > 1 and rank == 1 % world:
                    assert C.broadcast(comm_s, t, 1 % world) is True
        # regroup -> (divide) -> allgather = aggregate, three rounds on the same table: K2 + allgather rows
        t, slab = make(dt, seed=2); g_ = twin(t)
        for rep_ in range(3):
            c_before = calls["n"]
            assert C.regroup(comm_s, t, Partitioner(world)); C.regroup(comm_g, g_, Partitioner(world)); torch.cuda.synchronize()
            assert sorted(t.get_partition_ids()) == [p_ for p_ in range(P) if p_ % world == rank]
            if dt == torch.float64:
                same(t, g_)
            for tt in (t, g_):
                for p_ in tt.get_partitions():
                    p_.get().mul_(0.5)
            assert C.allgather(comm_s, t); C.allgather(comm_g, g_); torch.cuda.synchronize()
            assert t.is_dense() and t.get_num_partitions() == P
            if dt == torch.float64:
                same(t, g_)
        # rotate: ids travel with the rows; ring and an arbitrary permutation
        ids = [1000 * rank + i for i in range(5)]
        t, _ = make(dt, seed=3, rows=5, ids=ids); g_ = twin(t)
        perm = {r: (r + 1) % world for r in range(world)} if world < 3 else {r: (r * 3 + 1) % world if (world % 3) else (r + 2) % world for r in range(world)}
        if sorted(perm.values()) != list(range(world)):
            perm = {r: (r + 1) % world for r in range(world)}
        for rmap in (None, perm, None):
            l0 = _native.launches()
            assert C.rotate(comm_s, t, rmap); C.rotate(comm_g, g_, rmap); torch.cuda.synchronize()
            assert world == 1 or _native.launches() > l0
            same(t, g_)
        # push / pull (K18): local = all P rows on every rank, global = the rows this rank owns
        nown = (P + world - 1) // world
        for use_bcast in (False, True):
            lt, lslab = make(dt, seed=4)
            gs = ws.alloc(nown * W, dt).view(nown, W); gs.zero_()
            owned = [p_ for p_ in range(P) if p_ % world == rank]
            gt = Table.dense(1, DoubleArrPlus(), gs[:len(owned)], owned)
            lt2, gt2 = twin(lt), twin(gt)
            l0 = _native.launches()
            assert C.push(comm_s, lt, gt, Partitioner(world)); C.push(comm_g, lt2, gt2, Partitioner(world)); torch.cuda.synchronize()
            assert world == 1 or _native.launches() > l0
            if dt == torch.float64:
                same(gt, gt2)
            lslab.zero_()
            for p_ in lt2.get_partitions():
                p_.get().zero_()
            assert C.pull(comm_s, lt, gt, use_bcast); C.pull(comm_g, lt2, gt2, use_bcast); torch.cuda.synchronize()
            if dt == torch.float64:
                same(lt, lt2)
            ref = gs[:len(owned)].clone()
    # steady state of the dense routes does no host metadata exchange: repeat every op once more and count
    t, _ = make(torch.float64, seed=9)
    C.allreduce(comm_s, t); C.broadcast(comm_s, t, 0)
    n0 = calls["n"]
    for _ in range(3):
        C.allreduce(comm_s, t); C.broadcast(comm_s, t, 0)
        C.regroup(comm_s, t, Partitioner(world)); C.allgather(comm_s, t)
    torch.cuda.synchronize()
    assert calls["n"] == n0, (calls["n"], n0)
    C.Comm.all_gather_object = orig_ago
    # the allocator really frees: allocate / free in a loop far beyond the slab size
    ws.reset()
    for i in range(200):
        a_ = ws.alloc(4 << 20, torch.float32); b_ = ws.alloc(1 << 20, torch.float32)
        ws.free(a_); ws.free(b_)
    assert ws.bytes_in_use() == 0
    ws.check()
    if rank == 0:
        print("table collectives on NVLink kernels ok", flush=True)

    # ---- NVLS (multimem) data path, when the slab is multicast-bound (backend "vmm") ----
    if ws.has_nvls:
        ws.reset()
        for n in (4, 1000 * 4, (1 << 20) + 4, 1290000):
            t = ws.alloc(n, torch.float32)
            src = torch.rand(n, device=dev, generator=g)
            ref = src.clone(); dist.all_reduce(ref)
            t.copy_(src); ws.nvls_allreduce_(t); torch.cuda.synchronize()
            torch.testing.assert_close(t, ref, rtol=1e-5, atol=1e-5)
            t.fill_(float(rank)); ws.nvls_broadcast_(t, world - 1); torch.cuda.synchronize()
            assert torch.equal(t, torch.full_like(t, float(world - 1)))
        per = 4096
        t = ws.alloc(per * world, torch.float32); t.zero_(); t[rank * per:(rank + 1) * per] = float(rank + 1)
        ws.nvls_allgather_(t); torch.cuda.synchronize()
        assert torch.equal(t, torch.arange(1, world + 1, device=dev, dtype=torch.float32).repeat_interleave(per))
        if rank == 0:
            print("nvls multimem allreduce / broadcast / allgather ok", flush=True)

    # timing of the 5 MB allreduce vs NCCL
    ws.reset()
    n = 10000 * 129
    t = ws.alloc(n, torch.float32); u = torch.rand(n, device=dev)
    arms = [(lambda: ws.allreduce_sum_(t), "p2p"), (lambda: dist.all_reduce(u), "nccl")]
    if ws.has_nvls:
        arms.insert(0, (lambda: ws.nvls_allreduce_(t), "nvls"))
    for fn, name in arms:
        for _ in range(5):
            fn()
        torch.cuda.synchronize(); dist.barrier()
        e0 = torch.cuda.Event(enable_timing=True); e1_ = torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(20):
            fn()
        e1_.record(); torch.cuda.synchronize()
        ms = torch.tensor([e0.elapsed_time(e1_) / 20], device=dev); dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        if rank == 0:
            print(f"allreduce 5.16MB {name}: {ms.item()*1e3:.1f} us", flush=True)
    dist.barrier()
    if rank == 0:
        print("SYMM_OK", flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
