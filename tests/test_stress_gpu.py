"""Stress tests of the cross-GPU flag protocols (stand-in for racecheck):
random per-rank delays, thousands of rounds, counters that wrap around 2^32.  Needs 2 GPUs."""
import os

import pytest
import torch
import torch.multiprocessing as mp

from tests import _stress_workers as S

pytestmark = pytest.mark.gpu


def _ngpu():
    return torch.cuda.device_count() if torch.cuda.is_available() else 0


def _spawn(fn, world, port, kw, out_dir, timeout=300):
    ctx = mp.get_context("spawn")
    procs = [ctx.Process(target=fn, args=(r, world, port, kw, out_dir)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(timeout)
    hung = [p for p in procs if p.is_alive()]
    for p in hung:
        p.kill()
    assert not hung, "worker hung"
    assert [p.exitcode for p in procs] == [0] * world
    return [torch.load(os.path.join(out_dir, f"rank{r}.pt"), weights_only=False) for r in range(world)]


@pytest.mark.skipif(_ngpu() < 2, reason="needs 2 GPUs")
@pytest.mark.parametrize("rows,width", [(8, 320), (512, 1024)])     # 10 KiB kernel pushes / 2 MiB copy-engine pushes
def test_boundary_push_flag_ack_under_random_delays_and_counter_wrap(tmp_path, free_port, rows, width):
    rounds = 2000 if rows == 8 else 300
    res = _spawn(S.boundary_stress_worker, 2, free_port, dict(rounds=rounds, rows=rows, width=width), str(tmp_path))
    for r in res:
        assert r["bad"] == 0, f"{r['bad']} elements read before / after their round"
        assert r["got"] == r["expect"]
        assert r["cnt"] == (r["start"] + rounds) & 0xFFFFFFFF and r["cnt"] < r["start"], "counter did not wrap"


@pytest.mark.skipif(_ngpu() < 2, reason="needs 2 GPUs")
def test_dp_ll_allreduce_under_random_delays_and_epoch_wrap(tmp_path, free_port):
    rounds = 2000
    res = _spawn(S.dp_ll_stress_worker, 2, free_port, dict(rounds=rounds), str(tmp_path))
    for r in res:
        assert r["ok"], (r["p0"], r["e0"])
        assert r["zeroed"]
        assert r["epoch"] == ((1 << 32) - 7 + rounds) & 0xFFFFFFFF
    assert torch.equal(res[0]["flat"], res[1]["flat"])
