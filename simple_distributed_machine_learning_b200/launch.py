"""Local launcher: start one process per rank on this host with the reference's flags.

The reference is started by hand, one terminal per rank (readme:19:
``python3 distributed.py --rank=0 --world_size=2 --interface=eth0 --master_addr=... --master_port=2308``).
This does the same thing for all ranks of one host:

    python -m simple_distributed_machine_learning_b200.launch --nproc 2 -- --epochs 1 --schedule 1f1b

Every child gets ``--rank=R --world_size=N --interface=lo --master_addr=127.0.0.1 --master_port=P``
(overridable) and ``LOCAL_RANK=R`` (one GPU per rank on a GPU box); if a rank dies the others are
terminated and its exit code is returned -- the reference would hang (``rpc_timeout=0``, ref:167).
"""
from __future__ import annotations

import argparse
import os
import signal
import socket
import subprocess
import sys
import time
from typing import List, Optional, Sequence


def _free_port() -> int:
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def launch(nproc: int, child_args: Sequence[str], interface: str = "lo", master_addr: str = "127.0.0.1",
           master_port: Optional[int] = None, script: Optional[str] = None, poll_s: float = 0.2) -> int:
    port = master_port or _free_port()
    cmd0 = [sys.executable, script] if script else [sys.executable, "-m", "simple_distributed_machine_learning_b200.cli"]
    procs: List[subprocess.Popen] = []
    for r in range(nproc):
        env = dict(os.environ, LOCAL_RANK=str(r))
        args = [f"--rank={r}", f"--world_size={nproc}", f"--interface={interface}",
                f"--master_addr={master_addr}", f"--master_port={port}", *child_args]
        procs.append(subprocess.Popen(cmd0 + args, env=env))
    code = 0
    try:
        live = set(range(nproc))
        while live:
            for r in sorted(live):
                rc = procs[r].poll()
                if rc is None:
                    continue
                live.discard(r)
                if rc != 0 and code == 0:
                    code = rc
                    print(f"[launch] rank {r} exited with {rc}; stopping the other ranks", file=sys.stderr)
                    for o in live:
                        procs[o].send_signal(signal.SIGTERM)
            time.sleep(poll_s)
    except KeyboardInterrupt:
        code = 130
    finally:
        deadline = time.time() + 10.0
        for p in procs:
            if p.poll() is None:
                try:
                    p.wait(max(0.1, deadline - time.time()))
                except subprocess.TimeoutExpired:
                    p.kill()
    return code


def main(argv: Optional[Sequence[str]] = None) -> int:
    ap = argparse.ArgumentParser(description="start one process per rank on this host",
                                 formatter_class=argparse.ArgumentDefaultsHelpFormatter)
    ap.add_argument("--nproc", type=int, default=2, help="ranks (= world_size) to start")
    ap.add_argument("--interface", type=str, default="lo")
    ap.add_argument("--master_addr", type=str, default="127.0.0.1")
    ap.add_argument("--master_port", type=int, default=0, help="0 = pick a free port")
    ap.add_argument("--script", type=str, default=None,
                    help="run this script per rank instead of the package CLI (e.g. simple_distributed.py)")
    ap.add_argument("rest", nargs=argparse.REMAINDER, help="-- followed by the flags passed to every rank")
    ns = ap.parse_args(argv)
    rest = ns.rest[1:] if ns.rest[:1] == ["--"] else ns.rest
    return launch(ns.nproc, rest, ns.interface, ns.master_addr, ns.master_port or None, ns.script)


if __name__ == "__main__":
    sys.exit(main())
