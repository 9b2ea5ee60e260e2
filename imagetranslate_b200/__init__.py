"""B200-native (sm_100a) beam-search decode path for ImageTranslate models.

Drop-in for the reference's `seq_gen.BeamDecoder` (reference src/seq_gen.py:27-48): same constructor,
same `forward` keyword arguments, same return type.  All per-step arithmetic runs in hand-written CUDA
kernels behind the C ABI of include/itb.h; PyTorch only owns device memory and streams.
"""
from .seq_gen import BeamDecoder, get_outputs_until_eos  # noqa: F401

__all__ = ["BeamDecoder", "get_outputs_until_eos"]
