"""simple_distributed_machine_learning_b200 -- a B200-native (sm_100a) pipeline-parallel
training framework with the capabilities of maduc238/simple_distributed_machine_learning.

Layout
  models/    stage modules (the reference LeNet split, wide-MLP stages)
  ops/       hand-written CUDA kernels (ctypes C ABI) + autograd wrappers
  parallel/  topology, pipeline schedules, stage executor, DP/TP collectives
  utils/     config/CLI, synthetic data, checkpoints, logging, timing
"""
__version__ = "0.1.0"

import os as _os

# More hardware work queues: compute, NCCL send and NCCL receive streams should not alias
# onto one queue (a spinning receive kernel must never sit in front of unrelated work).
# Only effective if set before the CUDA context is created, hence at import time.
_os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")
