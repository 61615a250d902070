"""ctypes binding of the C ABI in include/mgard_b200.h.

The shared library is built in-tree by ``__graft_entry__.build()`` /
``mgard_b200/csrc/build.sh``.  There is NO fallback: if the library (or a CUDA
device) is missing every call raises.
"""
import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libmgard_b200.so")

c_u64 = ctypes.c_uint64
c_u32 = ctypes.c_uint32
c_dbl = ctypes.c_double
c_int = ctypes.c_int
c_vp = ctypes.c_void_p

# name -> (restype, argtypes); must list every symbol of include/mgard_b200.h
SIGNATURES = {
    "mgx_create": (c_int, [ctypes.POINTER(c_vp), c_int, c_vp]),
    "mgx_destroy": (None, [c_vp]),
    "mgx_last_error": (ctypes.c_char_p, [c_vp]),
    "mgx_version": (ctypes.c_char_p, []),
    "mgx_plan": (c_int, [c_vp, c_int, ctypes.POINTER(c_u64), c_int, c_u32,
                         c_u32, c_u64]),
    "mgx_levels": (c_int, [c_vp, ctypes.POINTER(c_int), ctypes.POINTER(c_u64)]),
    "mgx_max_compressed_bytes": (c_u64, [c_vp]),
    "mgx_norm_partial": (c_int, [c_vp, c_vp, c_int, ctypes.POINTER(c_dbl)]),
    "mgx_finish_norm": (c_dbl, [c_int, c_int, c_dbl, c_u64, c_int]),
    "mgx_decompose": (c_int, [c_vp, c_vp, c_int]),
    "mgx_recompose": (c_int, [c_vp, c_vp, c_int]),
    "mgx_quantize": (c_int, [c_vp, c_vp, c_int, c_dbl, c_dbl, c_dbl, c_vp]),
    "mgx_dequantize": (c_int, [c_vp, c_vp, c_int, c_dbl, c_dbl, c_dbl, c_vp]),
    "mgx_huff_encode": (c_int, [c_vp, c_vp, c_u64, c_vp, c_u64,
                                ctypes.POINTER(c_u64)]),
    "mgx_huff_decode": (c_int, [c_vp, c_vp, c_u64, c_vp, c_u64]),
    "mgx_huff_codebook": (c_int, [c_vp, c_vp, c_vp, c_vp,
                                  ctypes.POINTER(c_int)]),
    "mgx_compress": (c_int, [c_vp, c_vp, c_int, c_dbl, c_dbl,
                             ctypes.POINTER(c_dbl), c_vp, c_u64,
                             ctypes.POINTER(c_u64)]),
    "mgx_decompress": (c_int, [c_vp, c_vp, c_u64, c_int, c_dbl, c_dbl, c_dbl,
                               c_vp]),
    "mgx_hier_quantize": (c_int, [c_vp, c_vp, c_int, c_dbl, c_dbl, c_vp]),
    "mgx_compress_host": (c_int, [c_vp, c_vp, c_int, c_dbl, c_dbl,
                                  ctypes.POINTER(c_dbl), c_vp, c_u64,
                                  ctypes.POINTER(c_u64)]),
    "mgx_decompress_host": (c_int, [c_vp, c_vp, c_u64, c_int, c_dbl, c_dbl,
                                    c_dbl, c_vp]),
    "mgx_hierarchy_levels": (c_int, [c_int, ctypes.POINTER(c_u64), c_u64,
                                     ctypes.POINTER(c_int),
                                     ctypes.POINTER(c_u64)]),
    "mgx_max_dim_decomposition": (c_int, [c_int, ctypes.POINTER(c_u64), c_u64,
                                          ctypes.POINTER(c_int),
                                          ctypes.POINTER(c_u64),
                                          ctypes.POINTER(c_u64)]),
    "mgx_subdomain_shape": (c_int, [c_int, ctypes.POINTER(c_u64), c_int, c_u64,
                                    c_u64, ctypes.POINTER(c_u64),
                                    ctypes.POINTER(c_u64)]),
    "mgx_local_abs_tol": (c_dbl, [c_int, c_int, c_dbl, c_dbl, c_dbl, c_u64]),
    "mgx_launch_count": (c_u64, [c_vp]),
    "mgx_set_timing": (c_int, [c_vp, c_int]),
    "mgx_elapsed_ms": (c_int, [c_vp, c_int, c_int,
                               ctypes.POINTER(ctypes.c_float)]),
}

_lib = None


def load():
    """Load libmgard_b200.so; raises (never falls back) when it is missing."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                "mgard_b200: %s is missing -- run __graft_entry__.build() "
                "(there is no CPU fallback)" % LIB_PATH)
        lib = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(lib, name)  # AttributeError if a symbol is missing
            fn.restype = res
            fn.argtypes = args
        _lib = lib
    return _lib
