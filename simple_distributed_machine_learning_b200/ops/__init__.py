"""Hand-written sm_100a kernels (``csrc/``) behind a ctypes C ABI, plus autograd wrappers."""
