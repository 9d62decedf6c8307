"""Property tests (hypothesis) of the L1 data model against plain-Python models: a Table is "union of ids, combine
on collision" (SURVEY §9.1 item 1), the partitioner is |id % N| with the sign rule for negative ids (item 5), KV
tables are dictionaries with a combiner, and the wire format round-trips arbitrary arrays."""
import numpy as np
import torch
from hypothesis import given, settings, strategies as st

from harp_b200.io import serial as S
from harp_b200.keyval import Int2LongKVTable, TypeCombiner
from harp_b200.partition import ArrayCombiner, Operation, Partition, Partitioner, PartitionStatus, Table
from harp_b200.resource import DoubleArray, IntArray, LongArray

OPS = {Operation.SUM: lambda a, b: a + b, Operation.MAX: np.maximum, Operation.MIN: np.minimum,
       Operation.MULTIPLY: lambda a, b: a * b, Operation.MINUS: lambda a, b: a - b}


@settings(max_examples=60, deadline=None)
@given(st.sampled_from(list(OPS)), st.lists(st.tuples(st.integers(-20, 20), st.lists(st.integers(-50, 50), min_size=3, max_size=3)),
                                            min_size=0, max_size=25))
def test_table_is_union_of_ids_with_combine_on_collision(op, adds):
    table, model = Table(0, ArrayCombiner(op)), {}
    for pid, vals in adds:
        status = table.add_partition(Partition(pid, torch.tensor(vals, dtype=torch.int64)))
        if pid in model:
            assert status is PartitionStatus.COMBINED
            model[pid] = OPS[op](model[pid], np.array(vals))
        else:
            assert status is PartitionStatus.ADDED
            model[pid] = np.array(vals)
    assert sorted(table.get_partition_ids()) == sorted(model) and table.get_num_partitions() == len(model)
    for pid, want in model.items():
        assert table.get_partition(pid).get().tolist() == want.tolist()


@settings(max_examples=100, deadline=None)
@given(st.integers(1, 64), st.integers(-10 ** 6, 10 ** 6))
def test_partitioner_matches_the_java_remainder_rule(n, pid):
    # Java: id % N keeps the sign of id; negative results are negated (Partitioner.java:36-43)
    java_rem = abs(pid) % n * (1 if pid >= 0 else -1)
    assert Partitioner(n).get_worker_id(pid) == (java_rem if java_rem >= 0 else -java_rem)
    ids = torch.tensor([pid, -pid, pid + n])
    got = Partitioner(n).worker_ids(ids).tolist() if hasattr(Partitioner(n), "worker_ids") else None
    if got is not None:
        assert got == [Partitioner(n).get_worker_id(int(i)) for i in ids.tolist()]


@settings(max_examples=50, deadline=None)
@given(st.lists(st.tuples(st.integers(-30, 30), st.integers(-1000, 1000)), max_size=60))
def test_kv_table_is_a_dict_with_a_combiner(pairs):
    t, model = Int2LongKVTable(0, TypeCombiner("sum")), {}
    for k, v in pairs:
        t.add_key_val(k, v)
        model[k] = model.get(k, 0) + v
    assert t.num_keys() == len(model)
    for k, v in model.items():
        assert t.get_val(k) == v
    if model:
        keys = torch.tensor(sorted(model)); vals = torch.ones(len(model), dtype=torch.int64)
        t.add_batch(keys, vals)
        assert all(t.get_val(k) == model[k] + 1 for k in model)


@settings(max_examples=50, deadline=None)
@given(st.lists(st.tuples(st.integers(-2 ** 31, 2 ** 31 - 1), st.sampled_from(["d", "i", "l"]),
                          st.lists(st.integers(-2 ** 31, 2 ** 31 - 1), max_size=12)), max_size=8))
def test_partition_lists_round_trip_through_the_wire_format(parts):
    cls = {"d": (DoubleArray, torch.float64), "i": (IntArray, torch.int32), "l": (LongArray, torch.int64)}
    objs = [Partition(pid, cls[k][0].wrap(torch.tensor(vals, dtype=cls[k][1]))) for pid, k, vals in parts]
    data = S.encode_trans_list(objs)
    assert len(data) == S.num_trans_list_bytes(objs)
    back = S.decode_partition_list(data)
    assert [p.id() for p in back] == [p.id() for p in objs]
    for a, b in zip(back, objs):
        assert type(a.get()) is type(b.get()) and torch.equal(a.get().get(), b.get().get())


@settings(max_examples=30, deadline=None)
@given(st.text(max_size=20), st.text(max_size=20), st.integers(0, 1000))
def test_data_heads_round_trip_any_names(ctx, op, wid):
    d = S.Data(S.SIMPLE_LIST, ctx, wid, [IntArray.wrap(torch.tensor([1], dtype=torch.int32))], op)
    head, body = d.encode()
    e = S.Data.decode(head, body)
    assert (e.context_name, e.operation_name, e.worker_id) == (ctx, op, wid)


def test_cvb_alpha_newton_recovers_the_dirichlet_mle():
    """The update of the asymmetric topic prior (contrib LDAMapper.java:476-600) is the MLE of a Dirichlet from
    the statistics sum_d log(theta_dk): with the psi(gamma)-psi(sum gamma) terms replaced by log-samples of a
    known Dirichlet it must land on that Dirichlet, from a far-away start, and stay positive on the way."""
    import torch
    from harp_b200.apps.contrib import LDACVBMapper
    g = torch.Generator().manual_seed(3)
    true = torch.tensor([0.3, 1.5, 4.0, 0.8], dtype=torch.float64)
    D = 20000
    theta = \
        torch._sample_dirichlet(true.expand(D, 4).contiguous(), generator=g)
    suff = torch.log(theta.clamp_min(1e-300)).sum(0)
    est = LDACVBMapper.update_vector_alpha(torch.full((4,), 0.1, dtype=torch.float64), D, suff)
    assert bool((est > 0).all())
    assert float(((est - true) / true).abs().max()) < 0.05, est
    # a fixed point: the gradient D (psi(sum a) - psi(a)) + s vanishes there
    grad = D * (torch.digamma(est.sum()) - torch.digamma(est)) + suff
    assert float(grad.abs().max()) / D < 1e-4


def _random_graph_csr(n, m, seed):
    import numpy as np
    import torch
    rs = np.random.default_rng(seed)
    und = {(min(a, b), max(a, b)) for a, b in rs.integers(0, n, (m, 2)) if a != b}
    adj = {v: sorted({b for a, b in und if a == v} | {a for a, b in und if b == v}) for v in range(n)}
    ptr = torch.zeros(n + 1, dtype=torch.int64)
    ptr[1:] = torch.cumsum(torch.tensor([len(adj[v]) for v in range(n)]), 0)
    idx = torch.tensor([u for v in range(n) for u in adj[v]], dtype=torch.int32)
    return adj, ptr, idx


def test_colour_coding_dp_counts_colourful_embeddings_exactly():
    """The compressed colour-coding DP (ops/subgraph.py + apps/graph.py::colour_coding_pass) against brute force: for
    a FIXED colouring the DP total is exactly the number of injective, edge-preserving maps of the template whose
    images carry pairwise different colours (colorcount_HJ.java:861-925)."""
    import itertools
    import torch
    from harp_b200.apps.graph import SubTemplate, colour_coding_pass
    n = 11
    adj, ptr, idx = _random_graph_csr(n, 26, 4)
    for k, edges in ((3, [(0, 1), (1, 2)]), (4, [(0, 1), (0, 2), (0, 3)]), (5, [(0, 1), (1, 2), (1, 3), (3, 4)]),
                     (6, [(0, 1), (0, 2), (2, 3), (2, 4), (4, 5)])):
        adj_t = {i: [] for i in range(k)}
        for a, b in edges:
            adj_t[a].append(b); adj_t[b].append(a)
        tmpl = SubTemplate(0, frozenset(range(k)), adj_t)
        colours = torch.randint(0, k, (n,), generator=torch.Generator().manual_seed(k))
        got = float(colour_coding_pass(ptr, idx, colours, tmpl, k)[:, 0].sum())
        want = 0
        for phi in itertools.permutations(range(n), k):
            if len({int(colours[v]) for v in phi}) == k and all(phi[b] in adj[phi[a]] for a, b in edges):
                want += 1
        assert got == want, (k, got, want)
