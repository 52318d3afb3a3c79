"""Log lines.

The two formats of the reference are kept character for character
(/root/reference/simple_distributed.py:114-117 and :130-132); only rank 0 prints them
in the reference, here the rank that owns the loss (last stage, DP replica 0) does.
A JSON-lines sink adds throughput / timing for the benchmark harness.
"""
from __future__ import annotations

import json
import sys
import time
from typing import Optional, TextIO


def train_line(epoch: int, batch_idx: int, batch_len: int, dataset_len: int, num_batches: int,
               loss: float) -> str:
    # ref:115-117
    return 'Train Epoch: {} [{}/{} ({:.0f}%)]\tLoss: {:.6f}'.format(
        epoch, batch_idx * batch_len, dataset_len, 100. * batch_idx / num_batches, loss)


def test_block(test_loss: float, correct: int, dataset_len: int) -> str:
    # ref:130-132
    return '\nTest set: Average loss: {:.4f}, Accuracy: {}/{} ({:.0f}%)\n'.format(
        test_loss, correct, dataset_len, 100.0 * correct / dataset_len)


class JsonlSink:
    def __init__(self, path: Optional[str] = None, stream: Optional[TextIO] = None):
        self._f = open(path, "a") if path else stream

    def write(self, **record) -> None:
        if self._f is None:
            return
        record.setdefault("ts", time.time())
        self._f.write(json.dumps(record) + "\n")
        self._f.flush()

    def close(self) -> None:
        if self._f not in (None, sys.stdout, sys.stderr):
            self._f.close()


def trace_line(summary: dict) -> str:
    """One line per rank from ``utils.trace.StepTracer.summary()`` (no counterpart in the reference)."""
    means = " ".join(f"{k}={v:.3f}" for k, v in summary.get("mean_ms", {}).items())
    return ("[trace] rank {rank} stage {stage}: step {step_ms:.3f} ms, busy {busy_ms:.3f} ms, "
            "bubble {bubble:.1%} over {steps} steps ({clock}); mean op ms: {means}").format(means=means, **summary)
