"""On-disk formats, file dealing, host schedulers and dymoro."""
import random
import threading
import time

import numpy as np
import pytest

from harp_b200.dymoro import MPTask, RotationMapper, Scheduler, create_rotation_order
from harp_b200.io import datagen, formats as F, multifile
from harp_b200.sched import DynamicScheduler, StaticScheduler, Task


def test_kmeans_binary_points_are_big_endian_doubles(tmp_path):
    pts = np.arange(12, dtype=np.float64).reshape(3, 4) + 0.5
    p = tmp_path / "data_0"
    F.write_kmeans_points_bin(str(p), pts)
    raw = p.read_bytes()
    assert len(raw) == 96 and raw[:8] == np.array(0.5, dtype=">f8").tobytes()      # DataOutputStream.writeDouble
    assert np.array_equal(F.read_kmeans_points_bin(str(p), 4), pts)


def test_text_formats_roundtrip(tmp_path):
    rng = np.random.default_rng(0)
    c = rng.random((5, 3)) * 1000
    F.write_centroids_text(str(tmp_path / "c"), c)
    assert np.array_equal(F.read_centroids_text(str(tmp_path / "c"), 3), c)         # repr() round-trips doubles
    F.write_points_text(str(tmp_path / "p"), c)
    assert np.array_equal(F.read_points_text(str(tmp_path / "p"), 3), c)
    F.write_evaluation(str(tmp_path / "evaluation"), "MSE", 7.25)
    assert (tmp_path / "evaluation").read_text() == "MSE : 7.25\n"
    assert F.read_evaluation(str(tmp_path / "evaluation")) == ("MSE", 7.25)
    F.write_dense_csv(str(tmp_path / "x.csv"), c)
    assert np.array_equal(F.read_dense_csv(str(tmp_path / "x.csv")), c)
    F.write_coo(str(tmp_path / "r"), [1, 2], [3, 4], [0.5, 4.0])
    r, cc, v = F.read_coo([str(tmp_path / "r")])
    assert r.tolist() == [1, 2] and cc.tolist() == [3, 4] and v.tolist() == [0.5, 4.0]
    F.write_mf_model(str(tmp_path / "W-0"), [4, 9], c[:2])
    assert (tmp_path / "W-0").read_text().startswith("4 : ")
    ids, m = F.read_mf_model(str(tmp_path / "W-0"))
    assert ids.tolist() == [4, 9] and np.array_equal(m, c[:2])


def test_csr_libsvm_lda_graph_formats(tmp_path):
    F.write_csr(str(tmp_path / "a.csr"), [0, 2, 3], [0, 2, 1], [1.0, 2.0, 3.0])
    lines = (tmp_path / "a.csr").read_text().splitlines()
    assert lines[0] == "1,3,4" and lines[1] == "1,3,2"                                # 1-based DAAL CSR
    ro, ci, va = F.read_csr(str(tmp_path / "a.csr"))
    assert ro.tolist() == [0, 2, 3] and ci.tolist() == [0, 2, 1] and va.tolist() == [1.0, 2.0, 3.0]
    X = np.array([[0, 1.5, 0], [2.0, 0, 3.0]]); y = np.array([1, -1])
    F.write_libsvm(str(tmp_path / "s"), X, y)
    X2, y2 = F.read_libsvm(str(tmp_path / "s"), 3)
    assert np.array_equal(X, X2) and y2.tolist() == [1, -1]
    F.write_lda_docs(str(tmp_path / "d"), ["d0", "d1"], [np.array([3, 3, 7]), np.array([1])])
    names, docs = F.read_lda_docs([str(tmp_path / "d")])
    assert names == ["d0", "d1"] and docs[0].tolist() == [3, 3, 7]
    wt = np.array([[0, 2, 0], [5, 0, 1]])
    F.write_lda_word_model(str(tmp_path / "wm"), [10, 11], wt)
    assert (tmp_path / "wm").read_text().splitlines()[0] == "10 1:2"
    assert np.array_equal(F.read_lda_word_model(str(tmp_path / "wm"), 3)[11], wt[1])
    F.write_adjacency(str(tmp_path / "g"), {0: [1, 2], 1: [0]})
    assert (tmp_path / "g").read_text().splitlines()[0] == "0\t1,2,"
    assert F.read_adjacency(str(tmp_path / "g"))[0].tolist() == [1, 2]
    F.write_template(str(tmp_path / "t"), 3, [(0, 1), (1, 2)])
    assert F.read_template(str(tmp_path / "t")) == (3, [(0, 1), (1, 2)])
    F.write_rendezvous(str(tmp_path / "job_1"), ["h0", "h1"])
    assert (tmp_path / "job_1" / "nodes").read_text() == "#0\nh0\nh1\n"
    assert (tmp_path / "job_1" / "tasks").read_text() == "0\t0\n1\t1\n" and (tmp_path / "job_1" / "lock").exists()
    assert F.read_nodes_file(str(tmp_path / "job_1" / "nodes")) == {0: ["h0", "h1"]}


def test_reference_datasets_parse():
    """The reference's own sample files load (they are read-only under /root/reference)."""
    import os
    base = "/root/reference/datasets"
    if not os.path.isdir(base):
        pytest.skip("reference datasets not mounted")
    km = os.path.join(base, "daal_kmeans/densedistri")
    f = sorted(os.listdir(km))[0]
    m = F.read_dense_csv(os.path.join(km, f))
    assert m.ndim == 2 and m.shape[0] > 10 and m.shape[1] >= 2
    tp = os.path.join(base, "daal_subgraph/templates/u3-1.template")
    nv, edges = F.read_template(tp)
    assert nv == 3 and len(edges) == 2


def test_multifile_dealing_matches_reference_rule():
    files = [f"f{i:02d}" for i in range(11)]
    s = multifile.multi_file_splits(files, 4)              # avg 2, rest 3 -> 3,3,3,2 consecutive
    assert [len(x) for x in s] == [3, 3, 3, 2] and sum(s, []) == files
    assert multifile.shard_files(files, 3, 4) == files[9:]
    assert multifile.multi_file_splits(files[:2], 4) == [["f00"], ["f01"]]
    assert multifile.shard_files(files[:2], 3, 4) == []
    assert multifile.records_for(["a"]) == [("a", "a")]


def test_datagen_shapes(tmp_path):
    paths = datagen.generate_kmeans_data(1000, 7, 4, str(tmp_path / "data"), seed=1)
    assert len(paths) == 4 and sum(F.read_kmeans_points_bin(p, 7).shape[0] for p in paths) == 1000
    c = datagen.generate_centroids(5, 7, str(tmp_path / "c" / "init_centroids"))
    assert c.shape == (5, 7) and c.max() < 1000
    r, cc, v = datagen.synth_ratings(100, 50, 2000)
    assert r.max() < 100 and cc.max() < 50 and v.shape == (2000,)
    docs = datagen.synth_corpus(20, 50, 4, 30)
    assert len(docs) == 20 and max(int(d.max()) for d in docs) < 50


class _Sq(Task):
    def __init__(self):
        self.acc = 0

    def run(self, x):
        self.acc += x
        return x * x


def test_dynamic_scheduler_semantics():
    s = DynamicScheduler([_Sq() for _ in range(4)])
    s.submit_all(range(50))
    assert not s.has_output()                 # not running yet: inputs are queued, not counted
    s.start()
    outs = []
    while s.has_output():
        outs.append(s.wait_for_output())
    assert sorted(outs) == [i * i for i in range(50)]
    assert sum(t.acc for t in s.get_tasks()) == sum(range(50))     # task-private state
    s.pause_now(); s.submit(3); assert not s.has_output()
    s.clean_input_queue(); s.start(); assert not s.has_output()
    s.submit(4); assert s.wait_for_output() == 16 and not s.has_error()
    s.stop()


def test_static_scheduler_per_task_queues_and_submitter():
    class Fwd(Task):
        def run(self, x):
            if self.get_task_id() == 0 and x < 3:
                self.get_submitter().submit(1, x + 100)
            return (self.get_task_id(), x)
    st = StaticScheduler([Fwd(), Fwd()])
    st.start()
    st.submit(0, 1)
    assert st.wait_for_output(0) == (0, 1)
    deadline = time.time() + 5
    while not st.has_output(1) and time.time() < deadline:
        time.sleep(0.01)
    assert st.wait_for_output(1) == (1, 101)
    st.pause(); st.start(); st.submit(1, 9); assert st.wait_for_output(1) == (1, 9)
    st.stop()


def test_block_scheduler_exclusivity_and_timer():
    busy_r, busy_c, viol, lock = set(), set(), [], threading.Lock()

    class T(MPTask):
        def do_run(self, c, r):
            with lock:
                if r in busy_r or c[0] in busy_c:
                    viol.append((r, c))
                busy_r.add(r); busy_c.add(c[0])
            time.sleep(0.001)
            with lock:
                busy_r.discard(r); busy_c.discard(c[0])
            return 1
    sch = Scheduler(5, 5, list(range(5)), 5000, [T() for _ in range(3)], seed=3)
    sch.schedule([[j] for j in range(5)])
    assert sch.get_num_v_items_trained() == 25 and not viol          # whole grid, never sharing a row/col
    sch.set_timer(4)
    sch.schedule([[j] for j in range(5)], record=True)
    assert 0 < sch.get_num_v_items_trained() < 25 and not viol       # timer-bounded partial pass
    sch.stop()


def test_rotation_orders_visit_every_worker_once_per_row():
    N = 4
    orders = create_rotation_order(random.Random(7), 3, N)
    assert len(orders) == (2 * N - 1) * 3 and orders[:N] == list(range(N))     # row 0 = identity placement
    rm = RotationMapper(orders, N)
    pos = {i: orders[i] for i in range(N)}                                      # block -> worker
    for row in range(2):
        visited = {i: {pos[i]} for i in range(N)}
        for step in range(N):
            m = rm.next_map()
            assert sorted(m) == list(range(N)) and sorted(m.values()) == list(range(N))   # a permutation
            pos = {i: m[w] for i, w in pos.items()}
            if step < N - 1:
                for i in range(N):
                    visited[i].add(pos[i])
        assert all(len(v) == N for v in visited.values())
    assert RotationMapper(None, N).next_map() is None
