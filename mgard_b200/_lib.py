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



class MgxConfig(ctypes.Structure):
    """mgx_config of include/mgard_b200_x.h"""
    _fields_ = [("dev_type", c_int), ("dev_id", c_int),
                ("domain_decomposition", c_int), ("decomposition", c_int),
                ("lossless", c_int), ("huff_dict_size", c_u32),
                ("huff_block_size", c_u32), ("reorder", c_u32),
                ("normalize_coordinates", c_int),
                ("auto_pin_host_buffers", c_int),
                ("compress_with_dryrun", c_int), ("auto_cache_release", c_int),
                ("log_level", c_int), ("max_larget_level", c_u64),
                ("max_memory_footprint", c_u64), ("block_size", c_u64),
                ("zstd_compress_level", c_int), ("lz4_block_size", c_u32)]


class MgxMetadata(ctypes.Structure):
    """mgx_metadata of include/mgard_b200_x.h"""
    _fields_ = [("version", c_u32 * 3), ("file_version", c_u32 * 3),
                ("D", c_u32), ("shape", c_u64 * 5), ("nonuniform", c_int),
                ("dtype", c_int), ("mode", c_int), ("linf", c_int),
                ("s", c_dbl), ("norm", c_dbl), ("tol", c_dbl),
                ("decomposed", c_int), ("dd_type", c_int), ("dd_dim", c_u32),
                ("dd_size", c_u64), ("decomposition", c_int),
                ("l_target", c_u32), ("lossless", c_int),
                ("huff_dict_size", c_u32), ("huff_block_size", c_u32),
                ("reorder", c_u32), ("backend", c_int)]


_cfgp = ctypes.POINTER(MgxConfig)
_szp = ctypes.POINTER(ctypes.c_size_t)

# name -> (restype, argtypes); must list every symbol of include/mgard_b200.h
# and include/mgard_b200_x.h
SIGNATURES = {
    "mgx_config_default": (None, [_cfgp]),
    "mgx_x_compress": (c_int, [c_int, c_int, ctypes.POINTER(c_u64), c_dbl, c_dbl,
                               c_int, c_vp, ctypes.POINTER(c_vp), _szp, _cfgp,
                               c_int]),
    "mgx_x_compress_coords": (c_int, [c_int, c_int, ctypes.POINTER(c_u64), c_dbl, c_dbl,
                                      c_int, c_vp, ctypes.POINTER(c_vp), _szp,
                                      ctypes.POINTER(c_vp), _cfgp, c_int]),
    "mgx_metadata_serialize_coords": (c_int, [ctypes.POINTER(MgxMetadata), c_vp, c_vp,
                                              c_u64, ctypes.POINTER(c_u64)]),
    "mgx_metadata_coords": (c_int, [c_vp, c_u64, c_vp, c_u64, ctypes.POINTER(c_u64)]),
    "mgx_x_compress_multi": (c_int, [c_int, c_int, ctypes.POINTER(c_u64), c_dbl, c_dbl,
                                     c_int, c_vp, ctypes.POINTER(c_vp), _szp,
                                     ctypes.POINTER(c_int), c_int, _cfgp, c_int]),
    "mgx_x_decompress_multi": (c_int, [c_vp, ctypes.c_size_t, ctypes.POINTER(c_vp),
                                       ctypes.POINTER(c_u64), ctypes.POINTER(c_int),
                                       ctypes.POINTER(c_int), ctypes.POINTER(c_int),
                                       c_int, _cfgp, c_int]),
    "mgx_x_decompress": (c_int, [c_vp, ctypes.c_size_t, ctypes.POINTER(c_vp),
                                 ctypes.POINTER(c_u64), ctypes.POINTER(c_int),
                                 ctypes.POINTER(c_int), _cfgp, c_int]),
    "mgx_x_release_cache": (c_int, [_cfgp]),
    "mgx_x_pin_memory": (c_int, [c_vp, c_u64, _cfgp]),
    "mgx_x_check_memory_pinned": (c_int, [c_vp, _cfgp]),
    "mgx_x_unpin_memory": (c_int, [c_vp, _cfgp]),
    "mgx_x_last_error": (ctypes.c_char_p, []),
    "mgx_x_have_zstd": (c_int, []),
    "mgx_x_have_lz4": (c_int, []),
    "mgx_x_lz4_wrap": (c_int, [ctypes.c_void_p, c_u64, c_u64, ctypes.c_void_p, c_u64,
                               ctypes.POINTER(c_u64)]),
    "mgx_x_lz4_unwrap": (c_int, [ctypes.c_void_p, c_u64, ctypes.c_void_p, c_u64,
                                 ctypes.POINTER(c_u64)]),
    "mgx_x_launch_count": (c_u64, []),
    "mgx_metadata_serialize": (c_int, [ctypes.POINTER(MgxMetadata), c_vp, c_u64,
                                       ctypes.POINTER(c_u64)]),
    "mgx_metadata_parse": (c_int, [c_vp, c_u64, ctypes.POINTER(MgxMetadata),
                                   ctypes.POINTER(c_u64)]),
    "mgx_metadata_size": (c_int, [c_vp, c_u64, ctypes.POINTER(c_u64)]),
    "mgx_crc32": (c_u32, [c_vp, c_u64]),
    "mgx_estimate_footprint": (c_u64, [c_int, ctypes.POINTER(c_u64), c_int,
                                       c_dbl, c_int]),
    "mgx_max_dim_by_memory": (c_int, [c_int, ctypes.POINTER(c_u64), c_int, c_dbl,
                                      c_u64, ctypes.POINTER(c_int),
                                      ctypes.POINTER(c_int),
                                      ctypes.POINTER(c_u64),
                                      ctypes.POINTER(c_u64)]),
    "mgx_create": (c_int, [ctypes.POINTER(c_vp), c_int, c_vp]),
    "mgx_destroy": (None, [c_vp]),
    "mgx_last_error": (ctypes.c_char_p, [c_vp]),
    "mgx_version": (ctypes.c_char_p, []),
    "mgx_plan": (c_int, [c_vp, c_int, ctypes.POINTER(c_u64), c_int, c_u32,
                         c_u32, c_u64]),
    "mgx_plan_grid": (c_int, [c_vp, c_int, ctypes.POINTER(c_u64), c_int, c_u32,
                              c_u32, c_u64, ctypes.POINTER(c_vp), c_int]),
    "mgx_levels": (c_int, [c_vp, ctypes.POINTER(c_int), ctypes.POINTER(c_u64)]),
    "mgx_max_compressed_bytes": (c_u64, [c_vp]),
    "mgx_norm_partial": (c_int, [c_vp, c_vp, c_int, ctypes.POINTER(c_dbl)]),
    "mgx_norm_partial_noalloc": (c_int, [c_int, c_vp, c_vp, c_int, c_u64, c_int, c_vp,
                                         ctypes.c_char_p, c_u64]),
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
    "mgx_last_histogram": (c_int, [c_vp, c_vp]),
    "mgx_last_redo_count": (c_u64, [c_vp]),
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
