"""CLI contract of the reference (/root/reference/simple_distributed.py:139-165)."""
import os

import pytest

from simple_distributed_machine_learning_b200.utils import config as C


def test_reference_flags_names_types_defaults():
    p = C.build_parser()
    a = p.parse_args(["--rank", "0"])
    assert a.rank == 0 and isinstance(a.rank, int)
    assert a.world_size == 2 and isinstance(a.world_size, int)
    assert a.interface == "eth0"
    assert a.master_addr == "localhost"
    assert a.master_port == "29500" and isinstance(a.master_port, str)   # ref:154: a string
    a = p.parse_args(["--rank=1", "--world_size=2", "--interface=lo", "--master_addr=192.168.10.128",
                      "--master_port=2308"])   # the README's command line
    assert (a.rank, a.interface, a.master_addr, a.master_port) == (1, "lo", "192.168.10.128", "2308")


def test_rank_is_mandatory():
    with pytest.raises(AssertionError, match="Must provide rank"):
        C.parse_args([])


def test_reference_hyperparameters_are_defaults():
    cfg = C.parse_args(["--rank", "0"])
    assert (cfg.batch_size, cfg.epochs, cfg.lr, cfg.momentum, cfg.log_interval) == (60, 10, 0.1, 0.5, 10)
    assert (C.BATCH_SIZE, C.EPOCHS, C.LEARNING_RATE, C.MOMENTUM, C.LOG_INTERVAL) == (60, 10, 0.1, 0.5, 10)
    assert cfg.train_samples == 6000 and cfg.test_samples == 1000   # ref:91-92


def test_env_export(monkeypatch):
    for k in ("MASTER_ADDR", "MASTER_PORT", "GLOO_SOCKET_IFNAME", "TP_SOCKET_IFNAME", "NCCL_SOCKET_IFNAME"):
        monkeypatch.delenv(k, raising=False)
    cfg = C.parse_args(["--rank", "1", "--interface", "lo", "--master_addr", "127.0.0.1", "--master_port", "2308"])
    C.export_env(cfg)
    assert os.environ["MASTER_ADDR"] == "127.0.0.1"
    assert os.environ["MASTER_PORT"] == "2308"
    assert os.environ["GLOO_SOCKET_IFNAME"] == "lo"
    assert os.environ["TP_SOCKET_IFNAME"] == "lo"
    assert os.environ["NCCL_SOCKET_IFNAME"] == "lo"


def test_inconsistent_topology_rejected():
    # the reference lets world_size > 2 ranks idle (ref:177-184); we refuse to start
    with pytest.raises(ValueError):
        C.parse_args(["--rank", "0", "--world_size", "3", "--dp", "2"])
    with pytest.raises(ValueError):
        C.parse_args(["--rank", "0", "--world_size", "4"])  # lenet has 2 stages
    cfg = C.parse_args(["--rank", "0", "--world_size", "8", "--dp", "4"])
    assert cfg.resolved_pp() == 2
    with pytest.raises(ValueError):
        C.parse_args(["--rank", "5", "--world_size", "2"])
    with pytest.raises(ValueError):
        C.parse_args(["--rank", "0", "--schedule", "lockstep", "--microbatches", "4"])


def test_metrics_jsonl_sink(tmp_path):
    """--metrics_jsonl: machine-readable loss / throughput / test records next to the reference's log lines."""
    import io
    import json
    from simple_distributed_machine_learning_b200.train import Trainer
    from simple_distributed_machine_learning_b200.utils.config import TrainConfig
    path = tmp_path / "m.jsonl"
    cfg = TrainConfig(rank=0, world_size=1, device="cpu", epochs=1, train_samples=240, test_samples=60,
                      log_interval=2, metrics_jsonl=str(path))
    t = Trainer(cfg, out=io.StringIO())
    t.fit()
    t.close()
    recs = [json.loads(l) for l in path.read_text().splitlines()]
    kinds = [r["kind"] for r in recs]
    assert kinds.count("train") == 2 and kinds.count("epoch") == 1 and kinds.count("test") == 1
    ep = next(r for r in recs if r["kind"] == "epoch")
    assert ep["steps"] == 4 and ep["samples_per_s"] > 0
    te = next(r for r in recs if r["kind"] == "test")
    assert te["total"] == 60 and 0 <= te["correct"] <= 60
