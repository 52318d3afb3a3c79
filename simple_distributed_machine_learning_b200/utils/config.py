"""Configuration and command line.

The five process-topology flags are the reference's, verbatim in name, type and default
(/root/reference/simple_distributed.py:139-160): ``--rank`` (int, mandatory),
``--world_size`` (int, 2), ``--interface`` (str, "eth0"), ``--master_addr`` (str,
"localhost"), ``--master_port`` (**str**, "29500").  The reference's module constants
(ref:18-22: batch 60, 10 epochs, lr 0.1, momentum 0.5, log every 10 batches) become
optional flags whose defaults reproduce it.  Everything else (pipeline / data / tensor
parallel degrees, micro-batches, schedule, model family, dtype, checkpoints) is new.
"""
from __future__ import annotations

import argparse
import dataclasses
import os
from dataclasses import dataclass
from typing import Optional, Sequence

# Reference hyper-parameters (ref:18-22)
BATCH_SIZE = 60
EPOCHS = 10
LEARNING_RATE = 0.1
MOMENTUM = 0.5
LOG_INTERVAL = 10

SCHEDULES = ("lockstep", "gpipe", "1f1b", "zb-h1", "zb-h2")
MODELS = ("lenet", "mlp")
DTYPES = ("fp32", "bf16", "fp8")
OPTIMIZERS = ("sgd", "adam")
P2P_MODES = ("auto", "symm", "native", "torch")
EXECUTORS = ("auto", "graph", "native", "python")


@dataclass
class TrainConfig:
    # --- process topology (reference flags) ---
    rank: int = 0
    world_size: int = 2
    interface: str = "eth0"
    master_addr: str = "localhost"
    master_port: str = "29500"
    # --- reference hyper-parameters ---
    batch_size: int = BATCH_SIZE
    epochs: int = EPOCHS
    lr: float = LEARNING_RATE
    momentum: float = MOMENTUM
    log_interval: int = LOG_INTERVAL
    # --- parallelism ---
    pp: int = 0                # 0 = world_size / (dp * tp)
    dp: int = 1
    tp: int = 1
    microbatches: int = 1
    schedule: str = "1f1b"
    virtual_stages: int = 1    # model chunks per rank (> 1: interleaved zero-bubble schedule, mlp only, dp = tp = 1)
    # --- model ---
    model: str = "lenet"
    hidden: int = 8192
    layers: int = 8            # total Linear layers of the wide MLP (split evenly over pp)
    classes: int = 0           # 0 = hidden (mlp) / 10 (lenet)
    dtype: str = "fp32"
    optimizer: str = "sgd"
    weight_decay: float = 0.0
    dropout: bool = True
    ref_eval_dropout: bool = False   # reproduce ref:75 (F.dropout active at eval time)
    # --- runtime ---
    backend: str = "auto"      # gloo | nccl | auto (nccl when CUDA is available)
    device: str = "auto"       # cpu | cuda | auto
    seed: int = 1
    synthetic: bool = True
    data_root: str = "data"
    train_samples: int = 6000  # ref:91 first tenth of MNIST train
    test_samples: int = 1000   # ref:92 first tenth of MNIST test
    ckpt_dir: Optional[str] = None
    resume: bool = False
    save_every: int = 0        # epochs; 0 = only at the end when ckpt_dir is set
    timeout_s: float = 300.0   # watchdog for p2p / collectives (the reference never times out)
    max_steps: int = 0         # stop after this many train steps per epoch (0 = all)
    fused_collectives: bool = True  # in-kernel DP/TP collectives instead of NCCL calls
    use_cuda_graph: bool = True
    fused_lenet: bool = True        # reference model on csrc/lenet_fused.cu (GPU) instead of ATen autograd
    # stage-boundary transport: auto (GPUs: symm, CPUs: torch) | symm (push into peer-mapped slots,
    # csrc/boundary.cu) | native (own NCCL engine, csrc/p2p_engine.cu) | torch (torch.distributed p2p)
    p2p: str = "auto"
    # who walks the op list: auto (graph where the step is capturable, else python) | graph (the op
    # list is compiled once into ONE CUDA graph per rank, a step is one launch) | native
    # (csrc/executor.cu: C++ loop over per-op CUDA graphs + the NCCL engine) | python (eager loop)
    executor: str = "auto"
    trace_dir: Optional[str] = None  # per-op timeline (CUDA events + NVTX) -> Chrome-trace JSON per rank
    trace_steps: int = 32
    metrics_jsonl: Optional[str] = None   # JSON-lines metrics (loss at log steps, epoch throughput, test results)

    def resolved_pp(self) -> int:
        if self.pp:
            return self.pp
        denom = self.dp * self.tp
        if self.world_size % denom:
            raise ValueError(f"world_size {self.world_size} not divisible by dp*tp = {denom}")
        return self.world_size // denom

    def validate(self) -> "TrainConfig":
        pp = self.resolved_pp()
        if pp * self.dp * self.tp != self.world_size:
            # the reference lets extra ranks idle (ref:177-184); we reject instead
            raise ValueError(
                f"pp({pp}) * dp({self.dp}) * tp({self.tp}) != world_size({self.world_size})")
        if not (0 <= self.rank < self.world_size):
            raise ValueError(f"rank {self.rank} outside [0, {self.world_size})")
        if self.schedule not in SCHEDULES:
            raise ValueError(f"unknown schedule {self.schedule!r}")
        if self.schedule == "lockstep" and self.microbatches != 1:
            raise ValueError("lockstep (reference behaviour) runs exactly one micro-batch")
        if self.model not in MODELS:
            raise ValueError(f"unknown model {self.model!r}")
        if self.dtype not in DTYPES:
            raise ValueError(f"unknown dtype {self.dtype!r}")
        if self.model == "lenet" and pp > 2:
            raise ValueError("the reference LeNet has exactly two stages (pp must be 1 or 2)")
        if self.model == "mlp" and self.layers % pp:
            raise ValueError(f"layers ({self.layers}) must be divisible by pp ({pp})")
        if self.microbatches < 1:
            raise ValueError("microbatches must be >= 1")
        if self.virtual_stages < 1:
            raise ValueError("virtual_stages must be >= 1")
        if self.virtual_stages > 1:
            if self.model != "mlp" or self.dp != 1 or self.tp != 1:
                raise ValueError("--virtual_stages > 1 (interleaved schedule) needs --model mlp and dp = tp = 1")
            if self.layers % (pp * self.virtual_stages):
                raise ValueError(f"layers ({self.layers}) must be divisible by pp * virtual_stages "
                                 f"({pp} * {self.virtual_stages})")
        if self.p2p not in P2P_MODES or self.executor not in EXECUTORS:
            raise ValueError(f"p2p must be one of {P2P_MODES} and executor one of {EXECUTORS}")
        return self

    def asdict(self) -> dict:
        return dataclasses.asdict(self)


def build_parser() -> argparse.ArgumentParser:
    parser = argparse.ArgumentParser(
        description="Distributed Machine Learning",
        formatter_class=argparse.ArgumentDefaultsHelpFormatter,
    )
    # ---- the reference's five flags (ref:144-156), same names / types / defaults ----
    parser.add_argument("--rank", type=int, metavar="R", help="Number of rank")
    parser.add_argument("--world_size", type=int, default=2, metavar="N", help="Number of workers")
    parser.add_argument("--interface", type=str, default="eth0", metavar="I",
                        help="Interface that current device is listening on. It will default to eth0 if not provided.")
    parser.add_argument("--master_addr", type=str, default="localhost", metavar="MA",
                        help="Address of master, will default to localhost if not provided. "
                             "Master must be able to accept network traffic on the address + port.")
    parser.add_argument("--master_port", type=str, default="29500", metavar="MP",
                        help="Port that master is listening on, will default to 29500 if not provided. "
                             "Master must be able to accept network traffic on the host and port.")
    # ---- reference constants as flags ----
    parser.add_argument("--batch_size", type=int, default=BATCH_SIZE)
    parser.add_argument("--epochs", type=int, default=EPOCHS)
    parser.add_argument("--lr", type=float, default=LEARNING_RATE)
    parser.add_argument("--momentum", type=float, default=MOMENTUM)
    parser.add_argument("--log_interval", type=int, default=LOG_INTERVAL)
    # ---- new ----
    parser.add_argument("--pp", type=int, default=0, help="pipeline stages (0 = world_size/(dp*tp))")
    parser.add_argument("--dp", type=int, default=1, help="data-parallel replicas per stage")
    parser.add_argument("--tp", type=int, default=1, help="tensor-parallel ranks per stage")
    parser.add_argument("--microbatches", type=int, default=1)
    parser.add_argument("--schedule", choices=SCHEDULES, default="1f1b")
    parser.add_argument("--virtual_stages", type=int, default=1,
                        help="model chunks per rank; > 1 selects the interleaved zero-bubble schedule (mlp, "
                             "dp = tp = 1): chunk c of rank r is virtual stage c * pp + r, the ramp a rank waits "
                             "for shrinks by this factor")
    parser.add_argument("--model", choices=MODELS, default="lenet")
    parser.add_argument("--hidden", type=int, default=8192)
    parser.add_argument("--layers", type=int, default=8)
    parser.add_argument("--classes", type=int, default=0)
    parser.add_argument("--dtype", choices=DTYPES, default="fp32")
    parser.add_argument("--optimizer", choices=OPTIMIZERS, default="sgd")
    parser.add_argument("--weight_decay", type=float, default=0.0)
    parser.add_argument("--no_dropout", dest="dropout", action="store_false")
    parser.add_argument("--ref_eval_dropout", action="store_true",
                        help="keep stage-1 dropout active during evaluation like the reference does")
    parser.add_argument("--backend", choices=("auto", "gloo", "nccl"), default="auto")
    parser.add_argument("--device", choices=("auto", "cpu", "cuda"), default="auto")
    parser.add_argument("--seed", type=int, default=1)
    parser.add_argument("--real_data", dest="synthetic", action="store_false",
                        help="read MNIST idx files from --data_root instead of synthetic data")
    parser.add_argument("--data_root", type=str, default="data")
    parser.add_argument("--train_samples", type=int, default=6000)
    parser.add_argument("--test_samples", type=int, default=1000)
    parser.add_argument("--ckpt_dir", type=str, default=None)
    parser.add_argument("--resume", action="store_true")
    parser.add_argument("--save_every", type=int, default=0)
    parser.add_argument("--timeout_s", type=float, default=300.0)
    parser.add_argument("--max_steps", type=int, default=0)
    parser.add_argument("--no_fused_collectives", dest="fused_collectives", action="store_false")
    parser.add_argument("--no_cuda_graph", dest="use_cuda_graph", action="store_false")
    parser.add_argument("--no_fused_lenet", dest="fused_lenet", action="store_false")
    parser.add_argument("--p2p", choices=P2P_MODES, default=os.environ.get("SDML_P2P", "auto"),
                        help="stage-boundary transport: symm = push into the neighbour's peer-mapped slot + flag "
                             "(csrc/boundary.cu, GPUs of one NVLink domain), native = own NCCL engine "
                             "(csrc/p2p_engine.cu), torch = torch.distributed p2p; auto = symm on GPUs, torch on CPUs")
    parser.add_argument("--executor", choices=EXECUTORS, default=os.environ.get("SDML_EXECUTOR", "auto"),
                        help="who walks the op list: graph = compiled once into one CUDA graph per rank "
                             "(needs --p2p symm for more than one stage), native = C++ loop over per-op CUDA graphs "
                             "(csrc/executor.cu, with --p2p native), python = eager loop; auto = graph when capturable")
    parser.add_argument("--trace_dir", type=str, default=None,
                        help="write a per-op timeline (trace_rank<R>.json, Chrome-trace format) and the "
                             "measured pipeline bubble here; runs the step eagerly (no CUDA graph)")
    parser.add_argument("--trace_steps", type=int, default=32)
    parser.add_argument("--metrics_jsonl", type=str, default=None,
                        help="append machine-readable metrics here (one JSON object per line): the loss at "
                             "every log step, samples/s per epoch, test loss / accuracy; written by the rank that "
                             "owns the loss")
    return parser


def parse_args(argv: Optional[Sequence[str]] = None) -> TrainConfig:
    args = build_parser().parse_args(argv)
    assert args.rank is not None, "Must provide rank argument."  # ref:160
    names = {f.name for f in dataclasses.fields(TrainConfig)}
    cfg = TrainConfig(**{k: v for k, v in vars(args).items() if k in names})
    return cfg.validate()


def export_env(cfg: TrainConfig) -> None:
    """Rendezvous + NIC selection through the environment, as ref:162-165 does.

    ``TP_SOCKET_IFNAME`` is kept for parity even though no TensorPipe agent is started;
    ``NCCL_SOCKET_IFNAME`` is added so the NCCL bootstrap uses the same NIC (payload
    goes over NVLink).
    """
    os.environ["MASTER_ADDR"] = cfg.master_addr
    os.environ["MASTER_PORT"] = str(cfg.master_port)
    os.environ["GLOO_SOCKET_IFNAME"] = cfg.interface
    os.environ["TP_SOCKET_IFNAME"] = cfg.interface
    os.environ.setdefault("NCCL_SOCKET_IFNAME", cfg.interface)
