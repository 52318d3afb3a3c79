#!/usr/bin/env python
"""Benchmark harness (driver contract in the task statement).

    python bench.py --gpus N --steps K --warmup W                  # this framework
    python bench.py --impl reference --gpus N --steps K --warmup W # the unmodified reference

Headline metric (BASELINE.json): **train samples/sec, device-timed, max over ranks**.
Headline config (the only workload the reference itself can run, so both arms measure the
same thing): the reference's 2-stage LeNet on MNIST-shaped synthetic batches of 60, fp32,
SGD(lr 0.1, momentum 0.5) -- /root/reference/simple_distributed.py:18-22, 26-83, 100-113.
At N > 1 the HEADLINE uses the layout that maximises samples/s for this latency-bound model: N
data-parallel replicas, each running both stages as one CUDA graph, joined by the fused
in-kernel all-reduce + SGD (replica batch 60, "weak" scaling).

The same JSON line carries, as extra keys, EVERY named config of BASELINE.json that fits on N GPUs
(``default_extras``): ``lenet_pp`` (configs 2 / 4: the 2-stage pipeline x N/2 replicas, 8 micro-batches,
1F1B, peer-mapped stage boundary, one CUDA graph per rank), ``mlp`` (config 3: 8-layer MLP, hidden 8192,
bf16, N pipeline stages; ``mlp_mb16`` / ``mlp_mb32`` at 8 GPUs), ``mlp_fp8`` (same, MX-fp8 GEMMs),
``mlp_dp`` (N replicas, fused wgrad reduce-scatter), ``mlp_tp`` / ``mlp_tp_fp8`` (config 5 layout: N/2
stages x 2-way tensor parallel, hidden 16384).  Every entry has ``vs_baseline`` = this framework / the
same layout on plain PyTorch ops + NCCL + cuBLAS measured in the same process (``baseline``), the
measured and the simulated pipeline bubble, and the fused paths actually taken (``config.paths``).

Timing: W warm-up steps, then K steps.  ``value`` is device time: every step is
bracketed by its own CUDA-event pair on the compute stream with an L2 flush (256 MiB
write) between pairs, summed over K steps, max over ranks; a barrier +
``torch.cuda.synchronize()`` brackets the whole region.  ``e2e`` is wall-clock through the
public API ``Trainer.train_batch`` (H2D of the step's batch from pinned memory + step +
D2H read of the loss, every step), no flush, max over ranks.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))

METRIC = "train samples/sec, device-timed max over ranks"




def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", choices=("ours", "reference"), default="ours")
    ap.add_argument("--config", choices=("lenet", "mlp"), default="lenet",
                    help="headline workload (default: the reference's own model)")
    ap.add_argument("--no-extra", action="store_true", help="skip the secondary measurements")
    ap.add_argument("--no-baseline", action="store_true",
                    help="skip the in-run plain-PyTorch + NCCL + cuBLAS arm (vs_baseline stays null)")
    ap.add_argument("--no-bubble", action="store_true", help="skip the measured-bubble timeline steps")
    ap.add_argument("--extras", type=str, default="",
                    help="comma list of secondary configs (default: every named config that fits, see "
                         "default_extras): mlp (N pipeline stages, bf16), mlp_mb16 (same, 16 micro-batches), mlp_fp8 (same, MX-fp8 "
                         "forward/dgrad/wgrad GEMMs), lenet_pp (the named "
                         "2-stage pipeline x N/2 replicas, 8 micro-batches 1F1B), mlp_tp (N/2 stages x 2-way TP, "
                         "hidden 16384), mlp_tp_fp8 (same layout, MX-fp8 GEMMs + NCCL collectives), "
                         "mlp_dp (N replicas, fused wgrad reduce-scatter)")
    ap.add_argument("--baseline-ops", action="store_true",
                    help="baseline (b) of BASELINE.md: same engine, but ATen/cuDNN/cuBLAS ops + NCCL collectives "
                         "instead of this repo's kernels (also set SDML_GEMM_BACKEND=cublas for the MLP GEMMs)")
    ap.add_argument("--p2p", choices=("auto", "symm", "native", "torch"), default=os.environ.get("SDML_P2P", "auto"),
                    help="pipeline transport: symm (push into peer-mapped slots, csrc/boundary.cu; auto picks it "
                         "on GPUs), native (csrc/p2p_engine.cu, NCCL) or torch.distributed NCCL p2p")
    ap.add_argument("--executor", choices=("auto", "graph", "native", "python"),
                    default=os.environ.get("SDML_EXECUTOR", "auto"),
                    help="op-list executor: graph (one CUDA graph per rank and step; auto picks it when the step "
                         "is capturable), native (csrc/executor.cu, per-op CUDA graphs) or the Python loop")
    ap.add_argument("--microbatches", type=int, default=0, help="0 = the named config's value")
    ap.add_argument("--schedule", type=str, default="")
    ap.add_argument("--mlp-steps", type=int, default=20)
    ap.add_argument("--mlp-warmup", type=int, default=5)
    ap.add_argument("--mlp-microbatch", type=int, default=2048)
    ap.add_argument("--mlp-microbatches", type=int, default=8)
    ap.add_argument("--hidden", type=int, default=8192)
    ap.add_argument("--layers", type=int, default=8)
    return ap.parse_args()


# --------------------------------------------------------------------------- clocks
class ClockSampler:
    """Samples nvidia-smi clocks / throttle reasons while the timed region runs."""

    QUERY = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
             "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int = 0):
        self.gpu_index = gpu_index
        self.proc = None
        self.lines = []
        self._t = None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.QUERY}", "--format=csv,noheader,nounits", "-lms", "25",
                 "-i", str(self.gpu_index)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except Exception:
            self.proc = None
            return self

        def reader():
            for line in self.proc.stdout:
                self.lines.append((time.perf_counter(), line.strip()))

        self._t = threading.Thread(target=reader, daemon=True)
        self._t.start()
        return self

    def stop(self):
        if self.proc is not None:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=2)
            except Exception:
                self.proc.kill()

    def summary(self, t0: float, t1: float) -> dict:
        sm, mx, pw = [], [], []
        reasons = set()
        inside = [x for x in self.lines if t0 <= x[0] <= t1 + 0.05]
        if len(inside) < 2:   # region shorter than the sampling period: take the nearest samples
            inside = sorted(self.lines, key=lambda x: abs(x[0] - 0.5 * (t0 + t1)))[:4]
        for ts, line in inside:
            f = [x.strip() for x in line.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2])); pw.append(float(f[3]))
            except ValueError:
                continue
            for name, val in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"),
                                 f[5:9]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        return {"sm_mhz": statistics.median(sm), "sm_max_mhz": max(mx), "power_w_max": max(pw),
                "reasons": sorted(reasons), "samples": len(sm)}


# --------------------------------------------------------------------------- ours
def dist_env():
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    return rank, world


def layout_for(config: str, n: int, args):
    """(pp, dp, tp, microbatches, schedule) of a config at n GPUs."""
    if config == "lenet":
        # Headline: the layout that maximises samples/s.  The reference model (21 840 parameters,
        # batch 60) is launch-latency bound, so every GPU runs both stages as one CUDA graph and
        # the GPUs are data-parallel replicas joined by the fused in-kernel all-reduce + SGD.
        mb = args.microbatches or 1
        return 1, n, 1, mb, args.schedule or ("lockstep" if mb == 1 else "1f1b")
    if config == "lenet_pp":
        # BASELINE.json configs 2 / 4: the 2-stage pipeline (x N/2 replicas), 8 micro-batches, 1F1B
        pp = 2
        return pp, max(1, n // pp), 1, args.microbatches or 8, args.schedule or "1f1b"
    if config in ("mlp_tp", "mlp_tp_fp8"):
        tp = 2
        pp = max(1, n // tp)
        return pp, 1, tp, args.mlp_microbatches, args.schedule or ("gpipe" if pp == 1 else "zb-h2")
    if config == "mlp_dp":
        return 1, n, 1, args.mlp_microbatches, "gpipe"
    # "mlp" (bf16) and "mlp_fp8" (MX-fp8 forward / dgrad / wgrad GEMMs) share the pipeline layout:
    # N stages, global batch 8 x 2048 rows in 8 micro-batches (the named config); "mlp_mbK" is the
    # same step cut into K micro-batches (16 x 1024 rows, 32 x 512 rows: smaller bubble on deep pipelines)
    pp = n
    mb = int(config[len("mlp_mb"):]) if config.startswith("mlp_mb") else args.mlp_microbatches
    return pp, 1, 1, mb, args.schedule or ("gpipe" if n == 1 else "zb-h2")


def measure(trainer, batches, steps: int, warmup: int, global_batch: int, sampler: ClockSampler, e2e: bool = True):
    """Returns dict(device_ms_total, e2e_s_total, launches, clocks, h2d, d2h) for `steps` steps."""
    import torch
    import torch.distributed as dist
    from simple_distributed_machine_learning_b200.ops import _lib

    dev = trainer.device
    distributed = dist.is_initialized()
    ctrl = trainer.groups.ctrl if trainer.groups is not None else None

    def barrier():
        if distributed:
            dist.barrier(group=ctrl)

    def maxreduce(v: float) -> float:
        if not distributed:
            return v
        t = torch.tensor([v], dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX, group=ctrl)
        return float(t.item())

    nb = len(batches)
    # ---- inputs staged on the device for the device-timed region ----
    dev_batches = []
    for data, target in batches:
        d, t, n = trainer._replica_slice(data, target)
        x, tt = trainer._to_device(d, t)
        dev_batches.append((x, tt, n // trainer.topo.tp))
    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=dev)

    for i in range(warmup):
        x, tt, n = dev_batches[i % nb]
        trainer.train_step_device(x, tt, n)
    extra_warm = 0
    while trainer.graph_ok and dev_batches[0][2] not in trainer._graphs and extra_warm < 6:
        # late capture (three eager steps first): keep the capture itself out of the timed region
        x, tt, n = dev_batches[(warmup + extra_warm) % nb]
        trainer.train_step_device(x, tt, n)
        extra_warm += 1
    torch.cuda.synchronize(dev)
    trainer.check_peers()

    # ---- device-timed: per-step event pairs, L2 flushed between pairs ----
    starts = [torch.cuda.Event(enable_timing=True) for _ in range(steps)]
    ends = [torch.cuda.Event(enable_timing=True) for _ in range(steps)]
    graph_launches = sum(trainer.launches_per_graph.values()) if trainer.graph_ok else 0
    barrier()
    torch.cuda.synchronize(dev)
    launches0 = _lib.launch_count()
    t_wall0 = time.perf_counter()
    for i in range(steps):
        x, tt, n = dev_batches[(warmup + i) % nb]
        flush.zero_()
        starts[i].record()
        trainer.train_step_device(x, tt, n)
        ends[i].record()
    torch.cuda.synchronize(dev)
    barrier()
    t_wall1 = time.perf_counter()
    launches = _lib.launch_count() - launches0
    if trainer.graph_ok and trainer.launches_per_graph:
        launches += steps * max(trainer.launches_per_graph.values())
    device_ms = sum(s.elapsed_time(e) for s, e in zip(starts, ends))
    device_ms = maxreduce(device_ms)
    clocks = sampler.summary(t_wall0, t_wall1)
    if not e2e:
        if distributed:
            lt = torch.tensor([launches], dtype=torch.float64)
            dist.all_reduce(lt, group=ctrl)
            launches = int(lt.item())
        del flush
        return {"device_ms": device_ms, "e2e_s": None, "launches": launches, "clocks": clocks,
                "h2d": 0, "d2h": 0, "graph_launches_per_step": graph_launches}

    # ---- end to end through the public API: H2D + step + D2H of the loss, every step ----
    pinned = [(d.pin_memory(), t.pin_memory()) for d, t in batches]
    h2d = 0
    d2h = 0
    # Every step: H2D of that step's batch from pinned memory, the step, and a D2H copy of its loss
    # into pinned memory (train_batch(async_loss=True) queues it behind the step).  The host reads
    # step k's loss right after it has LAUNCHED step k+1 -- every one of the `steps` losses is read
    # inside the timed region, the last one before the clock stops -- so the read-back latency
    # overlaps the next step instead of idling the GPU (what a real training loop's logging does).
    for i in range(min(4, warmup)):
        loss = trainer.train_batch(*pinned[i % nb], async_loss=True)
        if loss is not None:
            float(loss)
    torch.cuda.synchronize(dev)
    barrier()
    torch.cuda.synchronize(dev)
    losses_read = 0
    t0 = time.perf_counter()
    prev = None
    for i in range(steps):
        data, target = pinned[(warmup + i) % nb]
        cur = trainer.train_batch(data, target, async_loss=True)
        if prev is not None:
            float(prev)                       # D2H read of the previous step's result
            losses_read += 1
        prev = cur
    if prev is not None:
        float(prev)
        losses_read += 1
    torch.cuda.synchronize(dev)
    barrier()
    e2e_s = maxreduce(time.perf_counter() - t0)
    assert losses_read in (0, steps), (losses_read, steps)   # 0 on ranks that own no loss
    d, t, n = trainer._replica_slice(*batches[0])
    if trainer.stage.is_first:
        h2d += d.numel() * d.element_size()
    if trainer.stage.is_last:
        h2d += t.numel() * t.element_size()
        d2h += 4
    if distributed:
        tt = torch.tensor([h2d, d2h], dtype=torch.float64)
        dist.all_reduce(tt, group=ctrl)
        h2d, d2h = int(tt[0].item()), int(tt[1].item())
        lt = torch.tensor([launches], dtype=torch.float64)
        dist.all_reduce(lt, group=ctrl)
        launches = int(lt.item())
    del flush
    return {"device_ms": device_ms, "e2e_s": e2e_s, "launches": launches, "clocks": clocks,
            "h2d": h2d, "d2h": d2h, "graph_launches_per_step": graph_launches}


def default_extras(world: int):
    """Secondary configs of the default run: every named config of BASELINE.json that fits on
    `world` GPUs (2: the 2-stage pipeline; 3: N-stage MLP; 4: 2 stages x N/2 replicas; 5: N/2 stages x
    2-way TP in bf16 and MX-fp8), plus the fp8 pipeline and the DP-only MLP."""
    if world == 1:
        return ["mlp", "mlp_fp8"]
    extras = ["lenet_pp", "mlp", "mlp_fp8", "mlp_dp", "mlp_tp", "mlp_tp_fp8"]
    if world >= 8:
        extras[2:2] = ["mlp_mb16", "mlp_mb32"]     # same step cut finer: smaller bubble on the deep pipeline
    return extras


def probe_fp8_peak(dev) -> dict:
    """Dense fp8 (E4M3, per-tensor scales) library GEMM 8192^3 through cuBLASLt, best of 10: the
    denominator of the fp8 roofline fractions (MEASURED_PEAKS.json only has the bf16 figure)."""
    import torch
    try:
        n = 8192
        a = torch.randn(n, n, device=dev).to(torch.float8_e4m3fn)
        b = torch.randn(n, n, device=dev).to(torch.float8_e4m3fn).t()
        one = torch.ones((), device=dev)
        for _ in range(3):
            torch._scaled_mm(a, b, scale_a=one, scale_b=one, out_dtype=torch.bfloat16)
        best = 1e9
        for _ in range(10):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            torch._scaled_mm(a, b, scale_a=one, scale_b=one, out_dtype=torch.bfloat16)
            e1.record()
            torch.cuda.synchronize(dev)
            best = min(best, e0.elapsed_time(e1))
        return {"fp8_tflops": 2.0 * n ** 3 / (best * 1e-3) / 1e12,
                "how": "torch._scaled_mm (cuBLASLt) E4M3 8192^3, per-tensor scales, best of 10, this process"}
    except Exception as e:  # noqa: BLE001 -- fall back to the nominal dense figure
        return {"fp8_tflops": 4500.0, "how": f"nominal dense fp8 peak (probe failed: {type(e).__name__})"}


def measure_bubble(trainer, batches, steps: int = 5) -> dict:
    """MEASURED pipeline bubble: a few extra steps run eagerly with a CUDA-event pair around every
    F / B / W / optimizer op (utils/trace.py); bubble = 1 - busy / step period, mean over ranks."""
    import torch
    import torch.distributed as dist
    from simple_distributed_machine_learning_b200.utils.trace import StepTracer
    tr = StepTracer(trainer.device, trainer.rank, trainer.pp_idx, skip=1, max_steps=steps, nvtx=False)
    trainer.engine.tracer = tr
    d, t, n = trainer._replica_slice(*batches[0])
    x, tt = trainer._to_device(d, t)
    for _ in range(steps + 2):
        trainer.engine.train_step(x, tt, n // trainer.topo.tp)
    torch.cuda.synchronize(trainer.device)
    trainer.engine.tracer = None
    sm = tr.summary()
    vals = torch.tensor([float(sm.get("bubble", 0.0)), float(sm.get("step_ms", 0.0))], dtype=torch.float64)
    if dist.is_initialized():
        dist.all_reduce(vals, group=trainer.groups.ctrl)
        vals /= dist.get_world_size(trainer.groups.ctrl)
    return {"bubble_pct_measured": round(100.0 * float(vals[0]), 2), "traced_step_ms": round(float(vals[1]), 3),
            "how": f"CUDA-event pair per op on the compute stream, {steps} eager steps, mean over ranks"}


def run_config(name: str, args, world: int, rank: int, sampler, baseline: bool, fp8_peak: dict) -> dict:
    import torch
    import torch.distributed as dist
    from simple_distributed_machine_learning_b200.models import mlp as mlp_mod
    from simple_distributed_machine_learning_b200.train import Trainer
    from simple_distributed_machine_learning_b200.utils.config import TrainConfig
    from simple_distributed_machine_learning_b200.utils.data import BatchLoader

    pp, dp, tp, mb, sched = layout_for(name, world, args)
    common = dict(rank=rank, world_size=world, interface=os.environ.get("GLOO_SOCKET_IFNAME", "lo"),
                  master_addr=os.environ.get("MASTER_ADDR", "127.0.0.1"),
                  master_port=os.environ.get("MASTER_PORT", "29500"),
                  device="cuda", pp=pp, dp=dp, tp=tp, microbatches=mb, schedule=sched, epochs=1,
                  p2p=args.p2p, executor=args.executor,
                  timeout_s=120.0)     # deadline of every host / in-kernel wait for a peer: a stuck config fails, the run goes on
    if baseline:
        # baseline (b) of BASELINE.md: the same engine and schedule, but ATen / cuDNN / cuBLAS ops,
        # NCCL collectives and NCCL p2p, launched eagerly from Python -- "plain PyTorch + NCCL + cuBLAS"
        common.update(fused_lenet=False, fused_collectives=False, use_cuda_graph=False, p2p="torch",
                      executor="python")
    mlp_mod._CUBLAS_BASELINE = bool(baseline)
    headline = name == args.config
    if name.startswith("lenet"):
        cfg = TrainConfig(model="lenet", batch_size=60, dtype="fp32", **common)
        steps, warmup = (args.steps, max(args.warmup, 3)) if headline else (40, 8)
    else:
        hidden = 2 * args.hidden if name.startswith("mlp_tp") else args.hidden
        gbatch = args.mlp_microbatch * args.mlp_microbatches   # rows per step, independent of mb
        dtype = "fp8" if (name.endswith("_fp8") and not baseline) else "bf16"
        cfg = TrainConfig(model="mlp", hidden=hidden, layers=args.layers, classes=hidden,
                          dtype=dtype, batch_size=gbatch // dp, train_samples=2 * gbatch,
                          test_samples=64, lr=0.01, momentum=0.5, **common)
        steps, warmup = (args.steps, max(args.warmup, 3)) if headline else (args.mlp_steps, max(args.mlp_warmup, 3))
    if baseline:
        steps, warmup = min(steps, 40), min(warmup, 8)
    trainer = Trainer(cfg, out=open(os.devnull, "w"))
    batches = list(BatchLoader(trainer.train_ds, cfg.batch_size * dp, drop_last=True))
    if name.startswith("mlp"):
        batches = [(d.to(torch.bfloat16), t) for d, t in batches]
    gb = cfg.batch_size * dp
    m = measure(trainer, batches, steps, warmup, gb, sampler, e2e=not baseline)
    value = gb * steps / (m["device_ms"] / 1e3)
    st = trainer.stage
    res = {
        "value": value, "ms_per_step": m["device_ms"] / steps, "steps": steps, "warmup": warmup,
        "gpu_launches": m["launches"], "clocks": m["clocks"],
        "config": {"global_batch": gb, "parallelism": trainer.topo.describe(), "microbatches": mb,
                   "schedule": sched, "optimizer": "sgd(lr,momentum=0.5)",
                   "l2": "256 MiB L2 flush between per-step CUDA-event pairs (device-timed); none for e2e",
                   "cuda_graph": bool(trainer.graph_ok), "p2p": trainer.engine.p2p_mode if pp > 1 else None,
                   "executor": ("graph" if trainer.graph_ok else ("native" if trainer._native_exec else "python")),
                   "baseline_ops": bool(baseline)},
    }
    if not baseline:
        res["e2e"] = {"value": gb * steps / m["e2e_s"], "unit": "samples/s", "h2d_bytes_per_step": m["h2d"],
                      "d2h_bytes_per_step": m["d2h"], "ms_per_step": 1e3 * m["e2e_s"] / steps}
        res["config"]["e2e_loss_read"] = "every step, asynchronous D2H into pinned memory, read one step late"
        # which of this repo's fused paths actually ran
        bnd = trainer.engine.bnd
        res["config"]["paths"] = {
            "fused_lenet_kernels": type(st).__name__ == "LeNetFusedStage",
            "fused_dp": (getattr(st, "_fused_dp", None) is not None or getattr(st, "_dp_w", None) is not None)
            if dp > 1 else None,
            # TP stages: bf16 = in-kernel AG->GEMM / GEMM->RS; fp8 = in-kernel GEMM->RS (+ opt-in push all-gather)
            "fused_tp": (getattr(st, "backend", None) == "fused" or bool(getattr(st, "fp8_fused_rs", False)))
            if tp > 1 else None,
            "tp_fp8_reduce_scatter_in_kernel": bool(getattr(st, "fp8_fused_rs", False)) if tp > 1 else None,
            "tp_fp8_all_gather_push": bool(getattr(st, "fp8_fused_ag", False)) if tp > 1 else None,
            "fp8_forward": bool(getattr(st, "fp8_forward", False) or getattr(st, "fp8", False)),
            "fp8_backward": bool(getattr(st, "fp8_backward", False) or getattr(st, "fp8", False)),
            "boundary_direct_gemm_writes": (bnd.direct_writes > 0) if bnd is not None else None,
        }
    if pp > 1:
        from simple_distributed_machine_learning_b200.parallel import schedule as _sched
        try:
            sim = _sched.simulate("gpipe" if sched == "lockstep" and mb > 1 else sched, pp, mb)
            res["bubble_pct_simulated"] = round(100.0 * sim.mean_bubble_fraction, 2)
        except ValueError:
            res["bubble_pct_simulated"] = None
        if not baseline and not args.no_bubble:
            try:
                res.update(measure_bubble(trainer, batches))
            except Exception as e:  # noqa: BLE001 -- the timing above stands on its own
                res["bubble_pct_measured"] = None
                res["bubble_error"] = f"{type(e).__name__}: {e}"[:200]
    if name.startswith("lenet"):
        res["config"].update({"model": "reference 2-stage LeNet (conv1,conv2 | fc1,fc2), MNIST-shaped 1x28x28",
                              "seq_len": None, "dtype": "fp32"})
    else:
        flops_local = st.flops_per_sample() * (gb // dp) * steps
        ft = torch.tensor([float(flops_local)], dtype=torch.float64)
        if dist.is_initialized():
            dist.all_reduce(ft, group=trainer.groups.ctrl)
        res["model_tflops"] = float(ft.item()) / (m["device_ms"] / 1e3) / 1e12
        peaks = ROOT / "MEASURED_PEAKS.json"
        try:
            pk = json.loads(peaks.read_text()) if peaks.exists() else {"bf16_tflops": 1649.3, "bf16_tflops_sustained": 1373.1}
            if cfg.dtype == "fp8":
                # fp8 runs are measured against a dense fp8 library GEMM probed in this process
                res["roofline"] = {"of_measured_fp8": res["model_tflops"] / (world * fp8_peak["fp8_tflops"]),
                                   "measured_fp8_tflops": fp8_peak["fp8_tflops"], "fp8_probe": fp8_peak["how"]}
            else:
                res["roofline"] = {
                    "of_measured_bf16_burst": res["model_tflops"] / (world * pk["bf16_tflops"]),
                    "of_measured_bf16_sustained": res["model_tflops"] / (world * pk["bf16_tflops_sustained"]),
                    "measured_bf16_tflops": pk["bf16_tflops"],
                    "measured_bf16_tflops_sustained": pk["bf16_tflops_sustained"]}
        except Exception:
            pass
        res["config"].update({"model": f"{args.layers}-layer MLP hidden {cfg.hidden} (Linear+bias+ReLU), "
                                       f"{cfg.hidden}-way softmax-CE",
                              "dtype": "mxfp8 (E4M3 + UE8M0/32) GEMMs, bf16 activations, fp32 master weights"
                              if cfg.dtype == "fp8" else "bf16",
                              "microbatch_rows": gbatch // dp // mb})
    trainer.close_local()
    del trainer
    mlp_mod._CUBLAS_BASELINE = False
    torch.cuda.empty_cache()
    return res


def run_ours(args) -> dict:
    import torch
    import torch.distributed as dist

    rank, world = dist_env()
    if world != args.gpus:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}: launch with torchrun --nproc-per-node {args.gpus}")
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a GPU (our arm)")
    local = int(os.environ.get("LOCAL_RANK", "0"))
    sampler = ClockSampler(local).start()
    results, baselines = {}, {}
    if args.no_extra:
        extras = []
    elif args.extras:
        extras = [e for e in args.extras.split(",") if e]
    else:
        extras = default_extras(world)
    extras = [e for e in extras if e != args.config]
    extras = [e for e in extras if not (e in ("lenet_pp", "mlp_tp", "mlp_tp_fp8", "mlp_dp") and world < 2)]
    configs = [args.config] + extras
    fp8_peak = {}
    try:
        if any(c.endswith("_fp8") for c in configs):
            torch.cuda.set_device(local)
            fp8_peak = probe_fp8_peak(torch.device("cuda", local))
        for name in configs:
            if args.baseline_ops:
                results[name] = run_config(name, args, world, rank, sampler, True, fp8_peak)
                continue
            try:
                results[name] = run_config(name, args, world, rank, sampler, False, fp8_peak)
            except Exception as e:  # noqa: BLE001 -- a failing secondary config must not lose the headline
                if name == args.config:
                    raise
                import traceback
                traceback.print_exc()
                results[name] = {"error": f"{type(e).__name__}: {e}"[:300]}
                continue
            if args.no_baseline:
                continue
            # in-run "plain PyTorch + NCCL + cuBLAS" arm of the same layout (fp8 variants are compared
            # with the bf16 baseline of their layout: there is no library fp8 training path to run)
            base_name = {"mlp_fp8": "mlp", "mlp_tp_fp8": "mlp_tp"}.get(name, name)
            if base_name not in baselines:
                baselines[base_name] = run_config(base_name, args, world, rank, sampler, True, fp8_peak)
            b = baselines[base_name]
            results[name]["vs_baseline"] = results[name]["value"] / b["value"]
            results[name]["baseline"] = {"value": b["value"], "ms_per_step": b["ms_per_step"], "steps": b["steps"],
                                         "config": base_name,
                                         "what": "same engine, schedule and layout on ATen/cuDNN/cuBLAS ops, NCCL "
                                                 "collectives and NCCL p2p, eager launches (bench.py --baseline-ops)"}
    finally:
        sampler.stop()
    head = results[args.config]
    out = {
        "metric": METRIC, "value": head["value"], "unit": "samples/s", "n_gpus": world,
        "steps": head["steps"], "warmup": head["warmup"], "ms_per_step": head["ms_per_step"],
        "higher_is_better": True, "scaling": "weak" if args.config == "lenet" else "strong",
        "vs_baseline": head.get("vs_baseline"), "dtype": head["config"]["dtype"], "data": "synthetic", "impl": "ours",
        "config": dict(head["config"],
                       vs_baseline_is="this framework / the same layout on plain PyTorch ops + NCCL + cuBLAS, "
                                      "measured in this run (BASELINE.md publishes no number)"),
        "clocks": head["clocks"], "e2e": head.get("e2e"), "gpu_launches": head["gpu_launches"],
    }
    if "baseline" in head:
        out["baseline"] = head["baseline"]
    for name in extras:
        if name in results:
            out[name] = results[name]
    if dist.is_initialized():
        dist.barrier()
        dist.destroy_process_group()
    return out if rank == 0 else {}


# --------------------------------------------------------------------------- reference
def run_reference(args) -> dict:
    """The unmodified reference script from baseline/_ref, through its own CLI, timed by
    baseline/run_reference.py.  It is a CPU program (its only CUDA touch moves the stage-1
    *input* to cuda:0 while the weights stay on the CPU, ref:63-67,71, which raises on a
    CUDA host), so its ranks run with CUDA hidden; extra ranks (N > 2) idle exactly as the
    reference makes them (ref:177-184)."""
    rank, world = dist_env()
    ref = ROOT / "baseline" / "_ref" / "simple_distributed.py"
    base = {"impl": "reference", "metric": METRIC, "unit": "samples/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "higher_is_better": True, "scaling": "weak",
            "dtype": "fp32", "data": "synthetic", "vs_baseline": None}
    if not ref.exists() and rank == 0 and Path("/root/reference").exists():
        # baseline/_ref is git-ignored: a fresh checkout has to (re)install it once
        subprocess.run([sys.executable, str(ROOT / "baseline" / "install_reference.py")],
                       stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    if not ref.exists():
        return {"impl": "reference", "unavailable": "baseline/_ref/simple_distributed.py missing (run baseline/install_reference.py)"} if rank == 0 else {}
    port = str(int(os.environ.get("MASTER_PORT", "29500")) + 17)
    runner = str(ROOT / "baseline" / "run_reference.py")
    env = dict(os.environ)
    env["CUDA_VISIBLE_DEVICES"] = ""
    env.pop("TORCHELASTIC_USE_AGENT_STORE", None)

    def launch(r, ws, threads, prt):
        return subprocess.Popen([sys.executable, runner, f"--rank={r}", f"--world_size={ws}",
                                 f"--master_port={prt}", f"--steps={args.steps}", f"--warmup={args.warmup}",
                                 f"--threads={threads}"],
                                stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, env=env)

    def parse(out):
        for line in out.splitlines():
            if line.startswith("REFBENCH "):
                return json.loads(line[len("REFBENCH "):])
        return None

    results = []
    if world == 1:
        # single launcher process: start the reference's two ranks ourselves
        ws = max(2, args.gpus)
        for k, threads in enumerate((0, 1)):
            prt = str(int(port) + k)
            procs = [launch(r, ws, threads, prt) for r in range(ws)]
            outs = []
            for p in procs:
                try:
                    outs.append(p.communicate(timeout=900)[0])
                except subprocess.TimeoutExpired:
                    p.kill()
                    outs.append("")
            r = parse(outs[0])
            if r:
                results.append(r)
    else:
        # under torchrun: rank i of the launcher becomes rank i of the reference
        for k, threads in enumerate((0, 1)):
            p = launch(rank, world, threads, str(int(port) + k))
            try:
                out = p.communicate(timeout=900)[0]
            except subprocess.TimeoutExpired:
                p.kill()
                out = ""
            r = parse(out) if rank == 0 else None
            if r:
                results.append(r)
    if rank != 0:
        return {}
    if not results:
        return {"impl": "reference", "unavailable": "reference run produced no timing (see baseline/run_reference.py)"}
    best = max(results, key=lambda r: r["samples_per_s"])
    base.update({
        "value": best["samples_per_s"], "ms_per_step": best["ms_per_step"],
        "e2e": {"value": best["samples_per_s"], "unit": "samples/s", "h2d_bytes_per_step": 0,
                "d2h_bytes_per_step": 0,
                "note": "the reference is a CPU program: host timer around ref:109-113, data loader included"},
        "gpu_launches": 0,
        "config": {"model": "reference 2-stage LeNet (Network1 | Network2), MNIST-shaped 1x28x28",
                   "global_batch": 60, "seq_len": None, "parallelism": "pp2 (TensorPipe RPC, lock-step)",
                   "device": "cpu", "threads_per_rank": best["threads"],
                   "runs": [{"threads": r["threads"], "samples_per_s": r["samples_per_s"]} for r in results],
                   "shim": "synthetic MNIST stand-in for torchvision download; rpc_timeout=0 -> 1e6 s "
                           "(torch>=2 treats 0 as 0 ms); CUDA hidden (script fails on a CUDA host)"},
        "timing": "host perf_counter around each dist_autograd.context block (no device work exists)",
    })
    return base


def main():
    args = parse_args()
    # The contract is ONE JSON line on stdout.  Libraries underneath write there too (NCCL
    # prints its version banner to stdout), so everything else is sent to stderr and the
    # JSON goes to the original stdout descriptor at the very end.
    sys.stdout.flush()
    real_stdout = os.dup(1)
    os.dup2(2, 1)
    out = run_reference(args) if args.impl == "reference" else run_ours(args)
    sys.stdout.flush()
    if out:
        os.write(real_stdout, (json.dumps(out) + "\n").encode())
    os.close(real_stdout)


if __name__ == "__main__":
    main()
