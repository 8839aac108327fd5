"""cpc2_b200 -- B200-native (sm_100a) implementation of the CPC2 training step.

Host side mirrors the reference's module interface (cpc/model.py, cpc/transformers.py,
cpc/criterion/criterion.py, cpc/train.py); all compute runs in hand-written CUDA kernels behind the
C ABI of libcpc_b200.so (include/cpc_b200.h).  There is no CPU or eager fallback.
"""
__version__ = "0.1.0"
