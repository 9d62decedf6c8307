"""Worker topology rules (reference unit tests ``worker/WorkerTest.java`` and ``NodesTest.java`` with the fixture
``src/test/resources/test_nodes``: two racks with one ``127.0.0.1`` each -> master 0, max 1, middle 0)."""
from harp_b200.topology import DEFAULT_WORKER_PORT_BASE, Nodes, Workers

FIXTURE = "#0\n127.0.0.1\n#1\n127.0.0.1\n"


def test_nodes_parsing_and_rack_order(tmp_path):
    p = tmp_path / "nodes"
    p.write_text(FIXTURE)
    n = Nodes.from_file(str(p))
    assert n.get_num_physical_nodes() == 2
    assert n.get_rack_list() == [0, 1] and n.get_nodes() == {0: ["127.0.0.1"], 1: ["127.0.0.1"]}
    assert n.get_node_list() == ["127.0.0.1", "127.0.0.1"]
    m = Nodes.from_lines(["#3", "a", "#1", "b", "c", "", "#3", "d"])
    assert m.get_rack_list() == [3, 1] and m.get_nodes()[3] == ["a", "d"]
    m.sort_racks()
    assert m.get_rack_list() == [3, 1]                       # 2 nodes each: stable
    k = Nodes.from_lines(["#0", "a", "#1", "b", "c"]); k.sort_racks()
    assert k.get_rack_list() == [1, 0]


def test_workers_ids_follow_the_reference_rules(tmp_path):
    p = tmp_path / "nodes"
    p.write_text(FIXTURE)
    w = Workers.from_file(str(p), 0)
    assert (w.get_master_id(), w.get_max_id(), w.get_middle_id(), w.get_min_id()) == (0, 1, 0, 0)
    assert w.is_master() and not w.is_max() and w.get_next_id() == 1 and w.get_num_workers() == 2
    assert w.get_self_info().port == DEFAULT_WORKER_PORT_BASE and w.get_next_info().rack == 1
    last = Workers.from_file(str(p), 1)
    assert last.is_max() and last.get_next_id() == 0 and not last.is_master() and last.is_self_in_worker()
    outsider = Workers.from_file(str(p))                     # not a member: next points at the master
    assert not outsider.is_self_in_worker() and outsider.get_next_id() == 0 and outsider.get_self_info() is None
    big = Workers(Nodes({0: ["h%d" % i for i in range(5)], 1: ["g0", "g1"]}), 3)
    assert big.get_max_id() == 6 and big.get_middle_id() == 3 and big.get_rack_workers() == {0: [0, 1, 2, 3, 4], 1: [5, 6]}
    assert [i.id for i in big.get_worker_info_list()] == list(range(7)) and big.get_worker_info(5).node == "g0"
    assert big.getMiddleID() == 3 and big.isTheOnlyWorker() is False


def test_workers_from_a_single_process():
    w = Workers.from_distributed()
    assert w.is_the_only_worker() and w.is_master() and w.get_next_id() == 0 and w.get_max_id() == 0
