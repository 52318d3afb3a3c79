"""simple_distributed_machine_learning_b200 -- a B200-native (sm_100a) pipeline-parallel
training framework with the capabilities of maduc238/simple_distributed_machine_learning.

Layout
  models/    stage modules (the reference LeNet split, wide-MLP stages)
  ops/       hand-written CUDA kernels (ctypes C ABI) + autograd wrappers
  parallel/  topology, pipeline schedules, stage executor, DP/TP collectives
  utils/     config/CLI, synthetic data, checkpoints, logging, timing
"""
__version__ = "0.1.0"
