"""B200-native GP-fitting hot path: drop-in for the numerics of rferguson22/Gaussian-Process
(GP_func.py, find_len_scales.py, calc_prob_surf.py).  Host code is Python/PyTorch (device memory,
streams, torch.distributed); every kernel lives in the hand-written CUDA library ``libgpk.so``
(C ABI in include/gpk.h).  There is no CPU fallback: importing the compute modules without the
built library, or calling them without a CUDA device, raises."""
__version__ = "0.1.0"
__all__ = ["engine", "pso", "surface", "lhs", "sharding", "_lib"]
