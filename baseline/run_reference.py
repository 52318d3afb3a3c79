"""Run the UNMODIFIED reference script (``baseline/_ref/simple_distributed.py``) as one rank,
through its own ``__main__`` / CLI, and time its training steps.

Two things outside the reference's control are substituted before ``runpy`` executes it:

* ``torchvision.datasets.MNIST`` -> a synthetic MNIST-shaped dataset (no network here;
  the reference calls ``MNIST(..., download=True)``, simple_distributed.py:87-88);
* ``rpc.TensorPipeRpcBackendOptions`` -> a subclass that turns ``rpc_timeout=0`` (meant as
  "infinite", ref:167) into a large finite value, because torch >= 2.x turns 0 into a
  0 ms store timeout and the rendezvous fails immediately.

Timing (rank 0 only): ``dist_autograd.context`` is wrapped so that enter->exit of each
``with`` block (= forward + loss + backward + optimizer step, ref:109-113) is one step.
After ``warmup + steps`` steps the run is stopped with a private exception and
``rpc.shutdown()`` is called (the reference's own tear-down, ref:186).  Result: one JSON
line on stdout prefixed with ``REFBENCH``.
"""
from __future__ import annotations

import argparse
import json
import os
import runpy
import sys
import time
from pathlib import Path

HERE = Path(__file__).resolve().parent
REF = HERE / "_ref" / "simple_distributed.py"


class _Done(Exception):
    pass


def main() -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--rank", type=int, required=True)
    ap.add_argument("--world_size", type=int, default=2)
    ap.add_argument("--master_port", type=str, default="29533")
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--threads", type=int, default=0, help="torch intra-op threads per rank (0 = default)")
    args = ap.parse_args()

    import torch
    import torch.distributed.autograd as dist_autograd
    import torch.distributed.rpc as rpc
    import torchvision
    from torch.utils.data import Dataset

    if args.threads:
        torch.set_num_threads(args.threads)
    os.environ.pop("TORCHELASTIC_USE_AGENT_STORE", None)   # we host our own store on --master_port

    class FakeMNIST(Dataset):
        """Same call shape as datasets.MNIST(root, train, download, transform) (ref:87-88)."""

        def __init__(self, root=None, train=True, download=False, transform=None, **_):
            sys.path.insert(0, str(HERE.parent))
            from simple_distributed_machine_learning_b200.utils.data import synthetic_mnist
            n = 60000 if train else 10000
            ds = synthetic_mnist(n, seed=1, train=train)
            self.images, self.labels = ds.images, ds.labels.tolist()

        def __len__(self):
            return len(self.labels)

        def __getitem__(self, i):
            return self.images[i], self.labels[i]

    torchvision.datasets.MNIST = FakeMNIST

    _Opts = rpc.TensorPipeRpcBackendOptions

    class PatchedOptions(_Opts):
        def __init__(self, *a, **kw):
            if kw.get("rpc_timeout", None) == 0:
                kw["rpc_timeout"] = 1.0e6
            super().__init__(*a, **kw)

    rpc.TensorPipeRpcBackendOptions = PatchedOptions

    times = []
    total = args.warmup + args.steps
    _orig_ctx = dist_autograd.context

    class TimedContext(_orig_ctx):
        def __enter__(self):
            self._t0 = time.perf_counter()
            return super().__enter__()

        def __exit__(self, et, ev, tb):
            r = super().__exit__(et, ev, tb)
            times.append(time.perf_counter() - self._t0)
            if len(times) >= total:
                raise _Done()
            return r

    if args.rank == 0:
        dist_autograd.context = TimedContext

    sys.argv = [str(REF), f"--rank={args.rank}", f"--world_size={args.world_size}", "--interface=lo",
                "--master_addr=127.0.0.1", f"--master_port={args.master_port}"]
    stopped_early = False
    try:
        runpy.run_path(str(REF), run_name="__main__")
    except _Done:
        stopped_early = True
        rpc.shutdown()
    if args.rank == 0:
        timed = times[args.warmup:args.warmup + args.steps]
        out = {"steps_timed": len(timed), "warmup": args.warmup, "total_s": sum(timed),
               "ms_per_step": 1e3 * sum(timed) / max(1, len(timed)), "batch": 60,
               "samples_per_s": 60.0 * len(timed) / max(1e-12, sum(timed)),
               "stopped_early": stopped_early, "cuda_visible": torch.cuda.is_available(),
               "threads": torch.get_num_threads()}
        print("REFBENCH " + json.dumps(out), flush=True)
    return 0


if __name__ == "__main__":
    sys.exit(main())
