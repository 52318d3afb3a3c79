"""Model stages: the reference's two-stage LeNet (/root/reference/simple_distributed.py:26-83) on ATen
and on fused CUDA kernels, and the wide-MLP stages of the scaling configs."""
