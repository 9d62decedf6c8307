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


def test_driver_turns_a_nodes_file_into_torchrun_commands(tmp_path, capsys):
    import pytest
    from harp_b200.launch import driver
    p = tmp_path / "nodes"
    p.write_text("#0\n127.0.0.1\n127.0.0.1\n#1\n127.0.0.1\n127.0.0.1\n")          # core/scripts/nodes: 4 local workers
    (host, cmd), = driver.plan(Nodes.from_file(str(p)), ["-m", "harp_b200.launch.kmeans", "1000"], 29611)
    assert host == "127.0.0.1" and "--nproc-per-node=4" in cmd and "--nnodes=1" in cmd
    assert cmd[cmd.index("--master-addr") + 1] == "127.0.0.1" and cmd[-3:] == ["-m", "harp_b200.launch.kmeans", "1000"]
    two = driver.plan(Nodes.from_lines(["#0", "n1", "n1", "#1", "n2", "n2"]), ["x.py"])
    assert [h for h, _ in two] == ["n1", "n2"]
    assert "--node-rank=1" in two[1][1] and "--rdzv-endpoint=n1:29500" in two[1][1] and "--nproc-per-node=2" in two[0][1]
    with pytest.raises(ValueError):
        driver.plan(Nodes.from_lines(["#0", "a", "a", "b"]), ["x.py"])
    assert driver.main(["--nodes", str(p), "--dry-run", "--", "-m", "harp_b200.launch.kmeans"]) == 0
    assert "torch.distributed.run" in capsys.readouterr().out


def test_driver_runs_local_workers(tmp_path):
    import subprocess, sys
    from pathlib import Path
    root = Path(__file__).resolve().parent.parent
    p = tmp_path / "nodes"
    p.write_text("#0\n127.0.0.1\n127.0.0.1\n")
    prog = tmp_path / "hello.py"
    prog.write_text("import os\nopen(os.path.join(os.path.dirname(__file__), 'r' + os.environ['RANK']), 'w').write(os.environ['WORLD_SIZE'])\n")
    res = subprocess.run([sys.executable, "-m", "harp_b200.launch.driver", "--nodes", str(p), "--master-port", "29612", "--", str(prog)],
                         cwd=str(root), capture_output=True, text=True, timeout=300)
    assert res.returncode == 0, res.stderr[-2000:]
    assert (tmp_path / "r0").read_text() == "2" and (tmp_path / "r1").read_text() == "2"
