"""Multi-GPU host logic: one MGARD-X subdomain per rank.

The reference treats subdomains as independent compress units (own hierarchy,
quantiser, codebook, container; no halo -- CompressionHighLevel/
GPUPipelines.hpp:89-205) coupled only through the global norm that a relative
bound needs (ErrorToleranceCalculator.hpp:45-52, 61-81) and the split of an L2
tolerance over subdomains (:126-147); afterwards every subdomain is compressed
with an ABS bound (CompressionHighLevel.hpp:126-129).  The only collective is
therefore one all-reduce of one fp64 scalar (MAX of |u|max for L-inf, SUM of
sum(u^2) otherwise) through torch.distributed: NCCL between GPUs, gloo in the
CPU tests.

Everything here is plumbing around the C ABI; no array arithmetic happens in
Python.
"""
import ctypes
import math
import struct

import numpy as np

from . import _lib

REL, ABS = 0, 1


def _u64(seq):
    return (ctypes.c_uint64 * len(seq))(*[int(x) for x in seq])


class DomainDecomposer:
    """MaxDim decomposition (DomainDecomposer.hpp:118-133, 192-223) with the
    stop rule 'at least one subdomain per device'."""

    def __init__(self, shape, num_devices):
        self.lib = _lib.load()
        self.shape = tuple(int(s) for s in shape)
        D = len(self.shape)
        dim, size, num = ctypes.c_int(), ctypes.c_uint64(), ctypes.c_uint64()
        rc = self.lib.mgx_max_dim_decomposition(
            D, _u64(self.shape), int(num_devices), ctypes.byref(dim),
            ctypes.byref(size), ctypes.byref(num))
        if rc != 0:
            from .compressor import MgardError
            raise MgardError(rc)
        self.dim, self.size, self.num_subdomains = dim.value, size.value, num.value

    def subdomain_shape(self, i):
        D = len(self.shape)
        out = (ctypes.c_uint64 * D)()
        off = ctypes.c_uint64()
        rc = self.lib.mgx_subdomain_shape(D, _u64(self.shape), self.dim, self.size,
                                          int(i), out, ctypes.byref(off))
        if rc != 0:
            raise IndexError("subdomain %d of %d" % (i, self.num_subdomains))
        return tuple(out), off.value

    def subdomain_slice(self, i):
        shp, off = self.subdomain_shape(i)
        sl = [slice(None)] * len(self.shape)
        sl[self.dim] = slice(off, off + shp[self.dim])
        return tuple(sl)

    def owner(self, i, world):
        """subdomain -> rank (round robin when there are more slabs than GPUs)"""
        return i % world


def frame(payloads):
    """body of the reference's stream: for each subdomain u64 size + payload
    (GPUPipelines.hpp:187-191)."""
    out = bytearray()
    for p in payloads:
        b = bytes(p)
        out += struct.pack("<Q", len(b))
        out += b
    return bytes(out)


def unframe(blob):
    out, pos = [], 0
    while pos < len(blob):
        (n,) = struct.unpack_from("<Q", blob, pos)
        pos += 8
        out.append(blob[pos:pos + n])
        pos += n
    return out


class ShardedCompressor:
    """One subdomain per rank.  ``comp`` is a Compressor planned for THIS rank's
    subdomain shape (anything with NormPartial / Compress / Decompress works:
    the CPU tests pass a stand-in).  ``group`` is a torch.distributed group or
    None (default group); with no initialised process group it degrades to a
    single subdomain."""

    def __init__(self, comp, dtype_code, total_elems, num_subdomains,
                 group=None, device=None):
        self.comp = comp
        self.dtype_code = dtype_code
        self.total_elems = int(total_elems)
        self.num_subdomains = int(num_subdomains)
        self.group = group
        self.device = device
        self.lib = _lib.load()

    # ---- the one collective ---------------------------------------------
    def all_reduce_partial(self, partial, linf):
        import torch
        import torch.distributed as dist
        if not (dist.is_available() and dist.is_initialized()):
            return float(partial)
        t = torch.tensor([partial], dtype=torch.float64,
                         device=self.device or "cpu")
        dist.all_reduce(t, op=dist.ReduceOp.MAX if linf else dist.ReduceOp.SUM,
                        group=self.group)
        return float(t.item())

    def global_norm(self, local, s):
        """calc_norm_decomposed (ErrorToleranceCalculator.hpp:61-123): per
        subdomain max|u| or sum u^2, combined, then sqrt( / N) for s != inf."""
        linf = math.isinf(s)
        part = self.all_reduce_partial(self.comp.NormPartial(local, s), linf)
        return float(self.lib.mgx_finish_norm(self.dtype_code, 1 if linf else 0,
                                              part, self.total_elems, 1))

    def local_abs_tol(self, ebtype, norm, tol, s):
        return float(self.lib.mgx_local_abs_tol(
            self.dtype_code, 1 if ebtype == REL else 0, norm, tol, s,
            self.num_subdomains))

    # ---- whole path --------------------------------------------------------
    def Compress(self, local, ebtype, tol, s, out=None):
        """returns (payload of this rank's subdomain, global norm, local tol)."""
        norm = self.global_norm(local, s) if ebtype == REL else 1.0
        ltol = self.local_abs_tol(ebtype, norm, tol, s)
        blob, _ = self.comp.Compress(local, ABS, ltol, s, 0.0, out=out)
        return blob, norm, ltol

    def Decompress(self, blob, ltol, s, out=None):
        return self.comp.Decompress(blob, ABS, ltol, s, 0.0, out=out)

    def gather_frames(self, payload_bytes, dst=0):
        """payloads of all ranks, framed in rank order, on rank dst (host side;
        the reference's output assembly, GPUPipelines.hpp:187-191)."""
        import torch.distributed as dist
        if not (dist.is_available() and dist.is_initialized()):
            return frame([payload_bytes])
        world = dist.get_world_size(self.group)
        rank = dist.get_rank(self.group)
        objs = [None] * world if rank == dst else None
        dist.gather_object(bytes(payload_bytes), objs, dst=dst, group=self.group)
        return frame(objs) if rank == dst else None
