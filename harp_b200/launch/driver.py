"""Start the workers listed in a nodes file — the role of ``collective/Driver.java`` + ``depl/Depl.java``.

The reference driver reads ``$HARP_HOME/scripts/nodes`` (``#rack`` lines, one host per worker), ``ssh``es to every
host and starts one JVM worker per line on port ``12800 + workerID``.  Here a worker is one process per GPU, started
by ``torchrun``; this driver turns the same nodes file into the torchrun command of every host (hosts that appear
k times get ``--nproc-per-node k``), runs it directly when all workers live on this machine, and prints (or, with
``--ssh``, executes over ssh) the per-host commands otherwise.

    python -m harp_b200.launch.driver --nodes nodes -- -m harp_b200.launch.kmeans 1000 10 100 5 2 2 10 /tmp/w /tmp/l
"""
from __future__ import annotations

import argparse
import shlex
import socket
import subprocess
import sys
from collections import OrderedDict
from typing import List, Sequence, Tuple

from ..topology import Nodes

LOCAL_NAMES = {"127.0.0.1", "localhost", "::1"}


def plan(nodes: Nodes, program: Sequence[str], master_port: int = 29500, python: str = sys.executable) -> List[Tuple[str, List[str]]]:
    """``[(host, argv)]`` — one torchrun invocation per distinct host, in first-appearance order; the first host is
    node rank 0 and the rendezvous endpoint (the master is worker 0, ``Workers.java:147``)."""
    per_host: "OrderedDict[str, int]" = OrderedDict()
    for h in nodes.get_node_list():
        per_host[h] = per_host.get(h, 0) + 1
    if not per_host:
        raise ValueError("nodes file lists no workers")
    counts = set(per_host.values())
    if len(counts) != 1:
        raise ValueError(f"torchrun needs the same number of workers on every host, nodes file has {dict(per_host)}")
    nproc = counts.pop()
    hosts = list(per_host)
    master = hosts[0]
    out = []
    for rank, h in enumerate(hosts):
        argv = [python, "-m", "torch.distributed.run", f"--nnodes={len(hosts)}", f"--nproc-per-node={nproc}"]
        if len(hosts) == 1:
            argv += ["--master-addr", "127.0.0.1" if master in LOCAL_NAMES else master, "--master-port", str(master_port)]
        else:
            argv += [f"--node-rank={rank}", "--rdzv-backend=c10d", f"--rdzv-endpoint={master}:{master_port}"]
        out.append((h, argv + list(program)))
    return out


def is_local(host: str) -> bool:
    return host in LOCAL_NAMES or host in (socket.gethostname(), socket.getfqdn())


def main(argv=None) -> int:
    ap = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawDescriptionHelpFormatter)
    ap.add_argument("--nodes", required=True, help="nodes file (#rack lines, one host per worker)")
    ap.add_argument("--master-port", type=int, default=29500)
    ap.add_argument("--dry-run", action="store_true", help="print the commands, start nothing")
    ap.add_argument("--ssh", action="store_true", help="start remote hosts' commands over ssh (like Depl.executeCMDandNoWait)")
    ap.add_argument("program", nargs=argparse.REMAINDER, help="-- followed by the torchrun program and its arguments")
    a = ap.parse_args(argv)
    program = a.program[1:] if a.program and a.program[0] == "--" else a.program
    if not program:
        ap.error("nothing to run: give the program after --")
    cmds = plan(Nodes.from_file(a.nodes), program, a.master_port)
    if a.dry_run:
        for host, cmd in cmds:
            print(f"[{host}] {shlex.join(cmd)}")
        return 0
    procs = []
    for host, cmd in cmds:
        if is_local(host):
            procs.append(subprocess.Popen(cmd))
        elif a.ssh:
            procs.append(subprocess.Popen(["ssh", host, shlex.join(cmd)]))
        else:
            print(f"[{host}] run there: {shlex.join(cmd)}", file=sys.stderr)
    rc = 0
    for p in procs:
        rc = p.wait() or rc
    return rc


if __name__ == "__main__":
    sys.exit(main())
