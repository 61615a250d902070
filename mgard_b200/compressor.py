"""Host-side mirror of the reference's low-level MGARD-X interface for the hot
path, on top of the C ABI (``include/mgard_b200.h``).

Names and argument meaning follow the reference:

* ``Config``      -- include/mgard-x/Config/Config.h:10-46 (fields the path uses)
* ``Hierarchy``   -- include/mgard-x/Hierarchy/Hierarchy.hpp:814-837
* ``Compressor``  -- include/mgard-x/CompressionLowLevel/Compressor.h:42-110
  (``Compress = CalculateNorm -> Decompose -> Quantize -> LosslessCompress ->
  Serialize``; ``Decompress`` the inverse)

PyTorch is used only to own device memory and streams.
"""
import ctypes
import math
from dataclasses import dataclass

import numpy as np
import torch

from . import _lib

REL, ABS = 0, 1  # mgard_x::error_bound_type (Utilities/Types.h:33)
INF = float("inf")

STATUS = {0: "Success", 1: "Failure", 2: "OutputTooLargeFailure",
          3: "NotSupportHigherNumberOfDimensionsFailure",
          4: "NotSupportDataTypeFailure", 5: "BackendNotAvailableFailure"}


class MgardError(RuntimeError):
    def __init__(self, status, msg=""):
        self.status = status
        super().__init__("%s%s" % (STATUS.get(status, str(status)),
                                   (": " + msg) if msg else ""))


@dataclass
class Config:
    """Subset of mgard_x::Config used by the hot path (defaults from
    src/mgard-x/Config/Config.cpp:32-65)."""
    dev_id: int = 0
    huff_dict_size: int = 8192
    huff_block_size: int = 1024
    normalize_coordinates: bool = True
    max_larget_level: int = 2 ** 64 - 1


def _dtype_code(dtype):
    if dtype in (torch.float32, np.float32, np.dtype("float32")):
        return 0
    if dtype in (torch.float64, np.float64, np.dtype("float64")):
        return 1
    raise MgardError(4, "only float32 / float64")


def infer_orthogonal_projection(D, s):
    """reference CompressionLowLevel/Compressor.h:38-40"""
    return not (math.isinf(s) and D <= 3)


class Hierarchy:
    """Level structure of a Cartesian grid (shape slowest->fastest): uniform
    (Hierarchy(shape, config), Hierarchy.hpp:814-837), or NON-UNIFORM when ``coords``
    gives one strictly increasing array of node coordinates per dimension
    (Hierarchy(shape, coords, config), Hierarchy.hpp:840-857)."""

    def __init__(self, shape, dtype=torch.float32, config=None, coords=None):
        self.shape = tuple(int(s) for s in shape)
        self.dtype = dtype
        self.config = config or Config()
        if len(self.shape) > 5:
            raise MgardError(3)
        self.dtype_code = _dtype_code(dtype)
        self.coords = None
        if coords is not None:
            npdt = np.float32 if self.dtype_code == 0 else np.float64
            if len(coords) != len(self.shape):
                raise MgardError(1, "one coordinate array per dimension")
            self.coords = [np.ascontiguousarray(c, npdt) for c in coords]
            for c, n in zip(self.coords, self.shape):
                if c.shape != (n,):
                    raise MgardError(1, "coordinate array length != dimension")

    def total_num_elems(self):
        return int(np.prod(self.shape))


class Compressor:
    """mgard_x::Compressor<D, T, CUDA> stand-in bound to one GPU stream."""

    def __init__(self, hierarchy, config=None, stream=None):
        self.lib = _lib.load()
        if not torch.cuda.is_available():
            raise MgardError(5, "no CUDA device (there is no CPU fallback)")
        self.hierarchy = hierarchy
        self.config = config or hierarchy.config
        self.device = torch.device("cuda", self.config.dev_id)
        self._ctx = ctypes.c_void_p()
        # without a caller stream the context runs on its own non-blocking
        # stream: inputs produced on torch's stream must be complete before a
        # call (the calls themselves are synchronous at return)
        self._own_stream = stream is None
        sptr = ctypes.c_void_p(stream.cuda_stream) if stream is not None else None
        self._check(self.lib.mgx_create(ctypes.byref(self._ctx),
                                        self.config.dev_id, sptr), ctx=False)
        self.Adapt(hierarchy, self.config)

    # ---- plumbing ---------------------------------------------------------
    def _check(self, rc, ctx=True):
        if rc != 0:
            msg = ""
            if ctx and self._ctx:
                msg = self.lib.mgx_last_error(self._ctx).decode()
            raise MgardError(rc, msg)

    def __del__(self):
        try:
            if getattr(self, "_ctx", None):
                self.lib.mgx_destroy(self._ctx)
                self._ctx = None
        except Exception:
            pass

    def Adapt(self, hierarchy, config=None):
        self.hierarchy = hierarchy
        self.config = config or self.config
        D = len(hierarchy.shape)
        shp = (ctypes.c_uint64 * D)(*hierarchy.shape)
        cptr = None
        if hierarchy.coords is not None:
            cptr = (ctypes.c_void_p * D)(*[c.ctypes.data for c in hierarchy.coords])
        self._check(self.lib.mgx_plan_grid(
            self._ctx, D, shp, hierarchy.dtype_code, self.config.huff_dict_size,
            self.config.huff_block_size, self.config.max_larget_level, cptr,
            1 if self.config.normalize_coordinates else 0))
        self.D = D
        self.N = hierarchy.total_num_elems()
        self.tdtype = (torch.float32 if hierarchy.dtype_code == 0
                       else torch.float64)

    def _sync_inputs(self):
        if self._own_stream:
            torch.cuda.current_stream(self.device).synchronize()

    def _dev(self, t, dtype=None):
        t = t.to(self.device)
        if dtype is not None and t.dtype != dtype:
            raise MgardError(4, "dtype mismatch")
        if not t.is_contiguous():
            t = t.contiguous()
        self._sync_inputs()
        return t

    def levels(self):
        L = ctypes.c_int()
        buf = (ctypes.c_uint64 * (40 * self.D))()
        self._check(self.lib.mgx_levels(self._ctx, ctypes.byref(L), buf))
        return L.value, [[buf[l * self.D + d] for d in range(self.D)]
                         for l in range(L.value + 1)]

    def max_compressed_bytes(self):
        return int(self.lib.mgx_max_compressed_bytes(self._ctx))

    def LastHistogram(self):
        """symbol histogram of the last Compress / DecomposeQuantizeFused /
        LosslessCompress call (numpy uint32[dict])."""
        h = np.empty(self.config.huff_dict_size, np.uint32)
        self._check(self.lib.mgx_last_histogram(self._ctx, h.ctypes.data))
        return h

    def LastRedoCount(self):
        """chains the partitioned Thomas solves of the last Compress redid sequentially"""
        return int(self.lib.mgx_last_redo_count(self._ctx))

    def launch_count(self):
        return int(self.lib.mgx_launch_count(self._ctx))

    TICKS = {"c_begin": 0, "c_norm": 1, "c_refactor": 2, "c_codebook": 3,
             "c_sizes": 4, "c_pack": 5, "c_end": 6, "d_begin": 8,
             "d_decode": 9, "d_dequant": 10, "d_end": 11}

    def set_timing(self, on=True):
        self._check(self.lib.mgx_set_timing(self._ctx, int(on)))

    def elapsed_ms(self, a, b):
        ms = ctypes.c_float()
        self._check(self.lib.mgx_elapsed_ms(self._ctx, self.TICKS[a],
                                            self.TICKS[b], ctypes.byref(ms)))
        return ms.value

    # ---- stages (Compressor.h:60-85) --------------------------------------
    def CalculateNorm(self, data, s):
        data = self._dev(data, self.tdtype)
        linf = 1 if math.isinf(s) else 0
        part = ctypes.c_double()
        self._check(self.lib.mgx_norm_partial(self._ctx, data.data_ptr(), linf,
                                              ctypes.byref(part)))
        return float(self.lib.mgx_finish_norm(
            self.hierarchy.dtype_code, linf, part.value, self.N,
            1 if self.config.normalize_coordinates else 0))

    def NormPartial(self, data, s):
        """local part of the norm (max |u| for s=inf, else sum u^2); combine
        partials of all subdomains (max / sum) and pass to FinishNorm."""
        data = self._dev(data, self.tdtype)
        part = ctypes.c_double()
        self._check(self.lib.mgx_norm_partial(
            self._ctx, data.data_ptr(), 1 if math.isinf(s) else 0,
            ctypes.byref(part)))
        return part.value

    def FinishNorm(self, partial, s, total_elems):
        return float(self.lib.mgx_finish_norm(
            self.hierarchy.dtype_code, 1 if math.isinf(s) else 0, partial,
            int(total_elems), 1 if self.config.normalize_coordinates else 0))

    def Decompose(self, data, orthogonal_projection):
        """in place on a device tensor (reordered level-blocked layout)."""
        assert data.is_cuda and data.is_contiguous() and data.dtype == self.tdtype
        self._sync_inputs()
        self._check(self.lib.mgx_decompose(self._ctx, data.data_ptr(),
                                           int(bool(orthogonal_projection))))
        return data

    def Recompose(self, data, orthogonal_projection):
        assert data.is_cuda and data.is_contiguous() and data.dtype == self.tdtype
        self._sync_inputs()
        self._check(self.lib.mgx_recompose(self._ctx, data.data_ptr(),
                                           int(bool(orthogonal_projection))))
        return data

    def Quantize(self, coeff, ebtype, tol, s, norm):
        coeff = self._dev(coeff, self.tdtype)
        codes = torch.empty(self.hierarchy.shape, dtype=torch.int64,
                            device=self.device)
        self._check(self.lib.mgx_quantize(self._ctx, coeff.data_ptr(),
                                          int(ebtype == REL), tol, s, norm,
                                          codes.data_ptr()))
        return codes

    def Dequantize(self, codes, ebtype, tol, s, norm):
        codes = self._dev(codes, torch.int64)
        out = torch.empty(self.hierarchy.shape, dtype=self.tdtype,
                          device=self.device)
        self._check(self.lib.mgx_dequantize(self._ctx, codes.data_ptr(),
                                            int(ebtype == REL), tol, s, norm,
                                            out.data_ptr()))
        return out

    def DecomposeQuantizeFused(self, data, ebtype, tol, norm):
        """hierarchical Decompose + Quantize in one pass (L-inf, D <= 3)."""
        data = self._dev(data, self.tdtype)
        codes = torch.empty(self.hierarchy.shape, dtype=torch.int64,
                            device=self.device)
        self._check(self.lib.mgx_hier_quantize(self._ctx, data.data_ptr(),
                                               int(ebtype == REL), tol, norm,
                                               codes.data_ptr()))
        return codes

    def HuffmanCodebook(self, freq):
        freq = self._dev(freq, torch.int32)  # uint32 bit pattern
        d = self.config.huff_dict_size
        cb = torch.empty(d, dtype=torch.int64, device=self.device)
        db = torch.empty(128 + d, dtype=torch.int64, device=self.device)
        ml = ctypes.c_int()
        self._check(self.lib.mgx_huff_codebook(self._ctx, freq.data_ptr(),
                                               cb.data_ptr(), db.data_ptr(),
                                               ctypes.byref(ml)))
        return cb, db, ml.value

    def LosslessCompress(self, codes, cap=None):
        codes = self._dev(codes, torch.int64).reshape(-1)
        cap = cap or self.max_compressed_bytes()
        out = torch.empty(cap, dtype=torch.uint8, device=self.device)
        nb = ctypes.c_uint64()
        self._check(self.lib.mgx_huff_encode(self._ctx, codes.data_ptr(),
                                             codes.numel(), out.data_ptr(),
                                             cap, ctypes.byref(nb)))
        return out[:nb.value]

    def LosslessDecompress(self, blob):
        blob = self._dev(blob, torch.uint8)
        codes = torch.empty(self.N, dtype=torch.int64, device=self.device)
        self._check(self.lib.mgx_huff_decode(self._ctx, blob.data_ptr(),
                                             blob.numel(), codes.data_ptr(),
                                             self.N))
        return codes

    # ---- whole path (Compressor.hpp:204-286) ---------------------------------
    def Compress(self, data, ebtype, tol, s, norm=0.0, out=None):
        """returns (compressed uint8 device tensor view, norm)."""
        data = self._dev(data, self.tdtype)
        if out is None:
            out = torch.empty(data.numel() * data.element_size() + (1 << 20),
                              dtype=torch.uint8, device=self.device)
        nb = ctypes.c_uint64()
        nrm = ctypes.c_double(norm)
        self._check(self.lib.mgx_compress(self._ctx, data.data_ptr(),
                                          int(ebtype == REL), tol, s,
                                          ctypes.byref(nrm), out.data_ptr(),
                                          out.numel(), ctypes.byref(nb)))
        return out[:nb.value], nrm.value

    def Decompress(self, blob, ebtype, tol, s, norm, out=None):
        blob = self._dev(blob, torch.uint8)
        if out is None:
            out = torch.empty(self.hierarchy.shape, dtype=self.tdtype,
                              device=self.device)
        self._check(self.lib.mgx_decompress(self._ctx, blob.data_ptr(),
                                            blob.numel(), int(ebtype == REL),
                                            tol, s, norm, out.data_ptr()))
        return out

    # host buffers (numpy), copies inside the call
    def CompressHost(self, array, ebtype, tol, s, norm=0.0, out=None):
        array = np.ascontiguousarray(array)
        if out is None:
            out = np.empty(array.nbytes + (1 << 20), np.uint8)
        nb = ctypes.c_uint64()
        nrm = ctypes.c_double(norm)
        self._check(self.lib.mgx_compress_host(
            self._ctx, array.ctypes.data, int(ebtype == REL), tol, s,
            ctypes.byref(nrm), out.ctypes.data, out.size, ctypes.byref(nb)))
        return out[:nb.value], nrm.value

    def DecompressHost(self, blob, ebtype, tol, s, norm, out=None):
        blob = np.ascontiguousarray(blob, np.uint8)
        if out is None:
            out = np.empty(self.hierarchy.shape,
                           np.float32 if self.hierarchy.dtype_code == 0
                           else np.float64)
        self._check(self.lib.mgx_decompress_host(
            self._ctx, blob.ctypes.data, blob.size, int(ebtype == REL), tol, s,
            norm, out.ctypes.data))
        return out
