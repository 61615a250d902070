"""mgard_b200 -- B200-native (sm_100a) implementation of the MGARD-X compression
hot path: multilevel decomposition / recomposition, linear quantisation and
Huffman coding behind the reference's low-level Compressor interface.

The compute lives in hand-written CUDA kernels (mgard_b200/csrc) exported through
a C ABI (include/mgard_b200.h); this package is the thin host-side mirror of the
reference interface.  No CPU fallback exists.
"""
from .compressor import (ABS, INF, REL, Compressor, Config, Hierarchy,
                         MgardError, infer_orthogonal_projection)
from ._lib import LIB_PATH, SIGNATURES, load

__all__ = ["ABS", "REL", "INF", "Compressor", "Config", "Hierarchy",
           "MgardError", "infer_orthogonal_projection", "LIB_PATH",
           "SIGNATURES", "load"]
