"""High-level API: the Python face of mgard_x::compress / mgard_x::decompress
(reference include/compress_x.hpp:31-178) over the C ABI of
include/mgard_b200_x.h.  Inputs may be numpy arrays (host pointers) or torch
CUDA tensors (device pointers); the output lives where the input lives, as in
the reference (CompressionHighLevel.hpp:141-154).  Plumbing only."""
import ctypes
import math

import numpy as np

from . import _lib
from .compressor import MgardError

REL, ABS = 0, 1          # error_bound_type (Utilities/Types.h:34)
AUTO, SERIAL, OPENMP, CUDA, HIP, SYCL = range(6)   # device_type
# lossless_type (Types.h:36-56): Huffman and Huffman_Zstd are built
HUFFMAN, HUFFMAN_LZ4, HUFFMAN_ZSTD, CPU_LOSSLESS = range(4)


def Config(**kw):
    """mgard_x::Config with the reference defaults (Config.cpp:31-66)."""
    cfg = _lib.MgxConfig()
    _lib.load().mgx_config_default(ctypes.byref(cfg))
    for k, v in kw.items():
        if not hasattr(cfg, k):
            raise AttributeError("Config has no field " + k)
        setattr(cfg, k, v)
    return cfg


def _is_tensor(x):
    return type(x).__module__.startswith("torch")


def _dtype_code(x):
    name = str(x.dtype).replace("torch.", "")
    if name == "float32":
        return 0
    if name == "float64":
        return 1
    return -1


def _check(rc):
    if rc != 0:
        raise MgardError(rc, _lib.load().mgx_x_last_error().decode())


def compress(data, tol, s, mode, config=None, out=None, coords=None):
    """-> compressed stream (numpy uint8 array for host input, torch uint8 CUDA
    tensor for device input).  ``out``: optional pre-allocated buffer of the
    same kind (output_pre_allocated = true).  ``coords``: one array of node
    coordinates per dimension -> the non-uniform overload (compress_x.hpp:74-100)."""
    lib = _lib.load()
    cfg = config if config is not None else Config()
    shape = tuple(int(v) for v in data.shape)
    D = len(shape)
    shp = (ctypes.c_uint64 * max(D, 1))(*shape)
    n = int(np.prod(shape)) if D else 0
    code = _dtype_code(data)
    es = 8 if code == 1 else 4
    on_dev = _is_tensor(data)
    if on_dev:
        import torch
        assert data.is_cuda and data.is_contiguous()
        if out is None:
            out = torch.empty(n * es + 1000000, dtype=torch.uint8,
                              device=data.device)
        in_ptr, out_ptr, cap = data.data_ptr(), out.data_ptr(), out.numel()
    else:
        data = np.ascontiguousarray(data)
        if out is None:
            out = np.empty(n * es + 1000000, dtype=np.uint8)
        in_ptr, out_ptr, cap = data.ctypes.data, out.ctypes.data, out.size
    p = ctypes.c_void_p(out_ptr)
    size = ctypes.c_size_t(cap)
    if coords is not None:
        npdt = np.float64 if code == 1 else np.float32
        arrs = [np.ascontiguousarray(c, npdt) for c in coords]
        if len(arrs) != D or any(a.shape != (n_,) for a, n_ in zip(arrs, shape)):
            raise MgardError(1, "one coordinate array of the dimension's length per "
                                "dimension")
        cptr = (ctypes.c_void_p * D)(*[a.ctypes.data for a in arrs])
        _check(lib.mgx_x_compress_coords(
            D, code, shp, float(tol), float(s), int(mode), ctypes.c_void_p(in_ptr),
            ctypes.byref(p), ctypes.byref(size), cptr, ctypes.byref(cfg), 1))
    else:
        _check(lib.mgx_x_compress(D, code, shp, float(tol), float(s), int(mode),
                                  ctypes.c_void_p(in_ptr), ctypes.byref(p),
                                  ctypes.byref(size), ctypes.byref(cfg), 1))
    return out[:size.value]


def compress_multi(data, tol, s, mode, devices, config=None):
    """mgx_x_compress_multi: host array -> "MGARD" stream (numpy uint8), the MaxDim
    subdomains compressed on the listed GPUs of this process, one host thread each."""
    lib = _lib.load()
    cfg = config if config is not None else Config()
    data = np.ascontiguousarray(data)
    shape = tuple(int(v) for v in data.shape)
    D = len(shape)
    shp = (ctypes.c_uint64 * max(D, 1))(*shape)
    code = _dtype_code(data)
    devs = (ctypes.c_int * len(devices))(*[int(d) for d in devices])
    p = ctypes.c_void_p()
    size = ctypes.c_size_t(0)
    _check(lib.mgx_x_compress_multi(D, code, shp, float(tol), float(s), int(mode),
                                    ctypes.c_void_p(data.ctypes.data), ctypes.byref(p),
                                    ctypes.byref(size), devs, len(devices),
                                    ctypes.byref(cfg), 0))
    try:
        return np.ctypeslib.as_array(
            ctypes.cast(p, ctypes.POINTER(ctypes.c_uint8)), (size.value,)).copy()
    finally:
        ctypes.CDLL(None).free(p)


def decompress_multi(stream, devices, config=None):
    """mgx_x_decompress_multi: host stream -> numpy array, the subdomain records
    decoded on the listed GPUs of this process, one host thread each."""
    lib = _lib.load()
    cfg = config if config is not None else Config()
    stream = np.ascontiguousarray(stream)
    md, _ = read_metadata(bytes(stream[:min(stream.size, 1 << 26)]))
    shape = tuple(int(md.shape[d]) for d in range(md.D))
    out = np.empty(shape, dtype=np.float64 if md.dtype == 1 else np.float32)
    devs = (ctypes.c_int * len(devices))(*[int(d) for d in devices])
    p = ctypes.c_void_p(out.ctypes.data)
    _check(lib.mgx_x_decompress_multi(ctypes.c_void_p(stream.ctypes.data), stream.size,
                                      ctypes.byref(p), None, None, None, devs,
                                      len(devices), ctypes.byref(cfg), 1))
    return out


def read_coords(stream):
    """node coordinates stored in a stream's header: list of float64 arrays, one per
    dimension, or None for a uniform grid"""
    lib = _lib.load()
    buf = np.frombuffer(bytes(stream), dtype=np.uint8)
    md, hb = read_metadata(buf[:min(buf.size, 1 << 26)])
    if not md.nonuniform:
        return None
    total = sum(int(md.shape[d]) for d in range(md.D))
    out = np.empty(total, np.float64)
    cnt = ctypes.c_uint64()
    _check(lib.mgx_metadata_coords(ctypes.c_void_p(buf.ctypes.data), buf.size,
                                   ctypes.c_void_p(out.ctypes.data), total,
                                   ctypes.byref(cnt)))
    res, at = [], 0
    for d in range(md.D):
        n_ = int(md.shape[d])
        res.append(out[at:at + n_].copy())
        at += n_
    return res


def read_metadata(stream_head):
    """parse the "MGARD" header from host bytes -> (MgxMetadata, header bytes)"""
    lib = _lib.load()
    buf = np.frombuffer(bytes(stream_head), dtype=np.uint8)
    md = _lib.MgxMetadata()
    nb = ctypes.c_uint64()
    rc = lib.mgx_metadata_parse(ctypes.c_void_p(buf.ctypes.data), buf.size,
                                ctypes.byref(md), ctypes.byref(nb))
    _check(rc)
    return md, nb.value


def decompress(stream, config=None, out=None):
    """-> array of the shape / dtype recorded in the stream, in the memory space
    of ``stream`` (``out``: optional pre-allocated array of that kind)."""
    lib = _lib.load()
    cfg = config if config is not None else Config()
    on_dev = _is_tensor(stream)
    if on_dev:
        import torch
        head = bytes(stream[:17].cpu().numpy())
    else:
        stream = np.ascontiguousarray(stream)
        head = bytes(stream[:17])
    total = ctypes.c_uint64()
    hb = np.frombuffer(head, dtype=np.uint8)
    _check(lib.mgx_metadata_size(ctypes.c_void_p(hb.ctypes.data), hb.size,
                                 ctypes.byref(total)))
    full = (bytes(stream[:total.value].cpu().numpy()) if on_dev
            else bytes(stream[:total.value]))
    md, _ = read_metadata(full)
    shape = tuple(md.shape[d] for d in range(md.D))
    if on_dev:
        if out is None:
            out = torch.empty(shape, dtype=torch.float64 if md.dtype == 1
                              else torch.float32, device=stream.device)
        in_ptr, out_ptr, nbytes = stream.data_ptr(), out.data_ptr(), stream.numel()
    else:
        if out is None:
            out = np.empty(shape, dtype=np.float64 if md.dtype == 1 else np.float32)
        in_ptr, out_ptr, nbytes = stream.ctypes.data, out.ctypes.data, stream.size
    p = ctypes.c_void_p(out_ptr)
    _check(lib.mgx_x_decompress(ctypes.c_void_p(in_ptr), nbytes, ctypes.byref(p),
                                None, None, None, ctypes.byref(cfg), 1))
    return out


def release_cache(config=None):
    cfg = config if config is not None else Config()
    _check(_lib.load().mgx_x_release_cache(ctypes.byref(cfg)))
