"""speq_b200 -- B200-native (sm_100a) implementation of SPeQ's scan hot path behind a C ABI (include/speq_cuda.h)."""
__version__ = "0.1.0"
