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


def bind_to_gpu_numa(dev_id):
    """Pin the calling process to the CPUs of the NUMA node the GPU hangs off, so that
    the pinned host buffers it allocates next (first touch) and the threads that feed
    the copies are local to the GPU's PCIe root: with 8 ranks per box the host-pointer
    path is otherwise limited by cross-socket traffic, not by PCIe.  Returns
    {"numa_node", "cpus"} or None when the topology is not exposed (then nothing is
    changed)."""
    import os
    try:
        import torch
        bus = torch.cuda.get_device_properties(dev_id).pci_bus_id
        dom = torch.cuda.get_device_properties(dev_id).pci_domain_id
        devn = torch.cuda.get_device_properties(dev_id).pci_device_id
        path = "/sys/bus/pci/devices/%04x:%02x:%02x.0/numa_node" % (dom, bus, devn)
        with open(path) as f:
            node = int(f.read().strip())
        if node < 0:
            return None
        with open("/sys/devices/system/node/node%d/cpulist" % node) as f:
            txt = f.read().strip()
        cpus = set()
        for part in txt.split(","):
            if "-" in part:
                a, b = part.split("-")
                cpus.update(range(int(a), int(b) + 1))
            elif part:
                cpus.add(int(part))
        allowed = os.sched_getaffinity(0)
        use = cpus & allowed
        if not use:
            return None
        os.sched_setaffinity(0, use)
        return {"numa_node": node, "cpus": len(use)}
    except Exception:
        return None


def serialize_header(shape, dtype_code, ebtype, tol, s, norm, dd_dim, dd_size,
                     decomposed=True, dict_size=8192, block_size=1024):
    """the "MGARD" header of a domain-decomposed stream (MetadataBase::Serialize,
    src/mgard-x/Metadata/Metadata.cpp:259-490; decomposition fields :353-372) through
    mgx_metadata_serialize -- same bytes mgx_x_compress writes for this domain."""
    lib = _lib.load()
    md = _lib.MgxMetadata()
    md.version[0], md.version[1], md.version[2] = 1, 0, 0
    md.D = len(shape)
    for d, n in enumerate(shape):
        md.shape[d] = int(n)
    md.dtype = dtype_code
    md.mode = 0 if ebtype == REL else 1
    md.linf = 1 if math.isinf(s) else 0
    if dtype_code == 0:
        # the templated reference entry takes tol, s and the norm as T
        # (CompressionHighLevel.hpp:311): what is recorded is their float value
        tol, norm = float(np.float32(tol)), float(np.float32(norm))
        s = s if math.isinf(s) else float(np.float32(s))
    md.s, md.norm, md.tol = s, norm, tol
    md.decomposed = 1 if decomposed else 0
    md.dd_type = 0
    md.dd_dim, md.dd_size = int(dd_dim), int(dd_size)
    md.decomposition = 0
    md.lossless = 0
    md.huff_dict_size, md.huff_block_size = dict_size, block_size
    md.backend = 3
    n = ctypes.c_uint64()
    lib.mgx_metadata_serialize(ctypes.byref(md), None, 0, ctypes.byref(n))
    buf = np.zeros(n.value, dtype=np.uint8)
    rc = lib.mgx_metadata_serialize(ctypes.byref(md), buf.ctypes.data, buf.size,
                                    ctypes.byref(n))
    if rc != 0:
        from .compressor import MgardError
        raise MgardError(rc, "mgx_metadata_serialize")
    return bytes(buf)


class ShardedDomain:
    """The subdomains ONE rank owns of a domain-decomposed field (or, with one
    shape per rank and no decomposition, its own field of a replicated run).

    ``shapes``: shapes of this rank's subdomains in subdomain order; one Compressor
    per distinct shape, all on ``stream``.  The only collective is the all-reduce of
    one fp64 scalar in ``global_norm`` (NCCL between GPUs, gloo in the CPU tests);
    ``group=False``: this object holds ALL subdomains, no collective at all;
    ``comp_factory`` lets the CPU tests pass a stand-in for the GPU Compressor."""

    def __init__(self, shapes, tdtype, total_elems, num_subdomains, group=None,
                 device=None, stream=None, dev_id=0, comp_factory=None):
        import torch
        self.lib = _lib.load()
        self.shapes = [tuple(int(x) for x in s) for s in shapes]
        self.tdtype = tdtype
        self.dtype_code = 0 if tdtype == torch.float32 else 1
        self.total_elems = int(total_elems)
        self.num_subdomains = int(num_subdomains)
        self.group, self.device, self.stream = group, device, stream
        self.comps = {}
        for shp in self.shapes:
            if shp not in self.comps:
                if comp_factory is not None:
                    self.comps[shp] = comp_factory(shp)
                else:
                    from .compressor import Compressor, Hierarchy, Config
                    self.comps[shp] = Compressor(
                        Hierarchy(shp, tdtype, Config(dev_id=dev_id)), stream=stream)
        self.last_norm_ms = 0.0
        self._timing = False
        self._stage = {}
        self._dev_in = None

    def comp(self, i):
        return self.comps[self.shapes[i]]

    def set_timing(self, on):
        self._timing = bool(on)
        for c in self.comps.values():
            c.set_timing(on)

    def launch_count(self):
        return sum(c.launch_count() for c in self.comps.values())

    def stage_ms(self, pairs):
        """stage times of the last Compress / Decompress summed over the subdomains"""
        return {k: sum(st.get(k, 0.0) for st in self._stage.values()) for k, _, _ in pairs}

    def _record(self, i, pairs):
        if self._timing:
            c = self.comp(i)
            self._stage[i] = {k: c.elapsed_ms(a, b) for k, a, b in pairs}

    # ---- the one collective ---------------------------------------------
    def all_reduce_partial(self, partial, linf):
        import torch
        import torch.distributed as dist
        if self.group is False or not (dist.is_available() and dist.is_initialized()):
            return float(partial)
        t = torch.tensor([partial], dtype=torch.float64, device=self.device or "cpu")
        dist.all_reduce(t, op=dist.ReduceOp.MAX if linf else dist.ReduceOp.SUM,
                        group=self.group)
        return float(t.item())

    def global_norm(self, fields, s):
        """calc_norm_decomposed (ErrorToleranceCalculator.hpp:45-52, 61-123): max |u|
        or sum u^2 per subdomain, combined over this rank's subdomains, all-reduced,
        then sqrt( / N) for s != inf."""
        import time
        linf = math.isinf(s)
        t0 = time.perf_counter()
        parts = [self.comp(i).NormPartial(f, s) for i, f in enumerate(fields)]
        self.last_norm_ms = (time.perf_counter() - t0) * 1e3
        # plain left-to-right fp64 adds like the reference's loop over subdomains
        # (ErrorToleranceCalculator.hpp:86-123); Python >= 3.12's sum() compensates
        local = 0.0
        for v in parts:
            local = max(local, v) if linf else local + v
        part = self.all_reduce_partial(local, linf)
        return float(self.lib.mgx_finish_norm(self.dtype_code, 1 if linf else 0, part,
                                              self.total_elems, 1))

    def local_abs_tol(self, ebtype, norm, tol, s):
        return float(self.lib.mgx_local_abs_tol(
            self.dtype_code, 1 if ebtype == REL else 0, norm, tol, s,
            self.num_subdomains))

    # ---- device-resident ---------------------------------------------------
    def Compress(self, fields, ebtype, tol, s, outs=None):
        """(payloads of this rank's subdomains, global norm, per-subdomain ABS bound)"""
        norm = self.global_norm(fields, s) if ebtype == REL else 1.0
        ltol = self.local_abs_tol(ebtype, norm, tol, s)
        blobs = []
        pairs = [("norm", "c_begin", "c_norm"), ("refactor_quantize", "c_norm", "c_refactor"),
                 ("codebook", "c_refactor", "c_codebook"), ("encode", "c_codebook", "c_sizes"),
                 ("pack", "c_sizes", "c_pack"), ("outliers_tail", "c_pack", "c_end")]
        for i, f in enumerate(fields):
            blob, _ = self.comp(i).Compress(f, ABS, ltol, s, 0.0,
                                            out=None if outs is None else outs[i])
            self._record(i, pairs)
            blobs.append(blob)
        return blobs, norm, ltol

    def Decompress(self, blobs, ltol, s, outs=None):
        res = []
        pairs = [("decode", "d_begin", "d_decode"), ("dequantize", "d_decode", "d_dequant"),
                 ("recompose", "d_dequant", "d_end")]
        for i, b in enumerate(blobs):
            res.append(self.comp(i).Decompress(b, ABS, ltol, s, 0.0,
                                               out=None if outs is None else outs[i]))
            self._record(i, pairs)
        return res

    # ---- host buffers (pinned numpy arrays): every byte crosses PCIe once --------
    def _staging(self):
        import torch
        if self._dev_in is None:
            self._dev_in = [torch.empty(s, dtype=self.tdtype, device=self.device)
                            for s in self.shapes]
            es = 4 if self.dtype_code == 0 else 8
            self._dev_out = [torch.empty(int(np.prod(s)) * es // 2 + (8 << 20),
                                         dtype=torch.uint8, device=self.device)
                             for s in self.shapes]
        return self._dev_in, self._dev_out

    def CompressHost(self, arrays, ebtype, tol, s, outs):
        """H2D of every subdomain, norm partials on the copies, all-reduce, compress
        from the resident copies, D2H of the payloads; returns numpy views of outs"""
        import torch
        dev_in, dev_out = self._staging()
        st = self.stream or torch.cuda.current_stream(self.device)
        with torch.cuda.stream(st):
            for d, a in zip(dev_in, arrays):
                d.copy_(torch.from_numpy(a), non_blocking=True)
        st.synchronize()
        blobs, norm, ltol = self.Compress(dev_in, ebtype, tol, s, outs=dev_out)
        res = []
        with torch.cuda.stream(st):
            for b, o in zip(blobs, outs):
                torch.from_numpy(o[:b.numel()]).copy_(b, non_blocking=True)
                res.append(o[:b.numel()])
        st.synchronize()
        return res

    def DecompressHost(self, hblobs, ltol, s, outs):
        import torch
        dev_in, dev_out = self._staging()
        st = self.stream or torch.cuda.current_stream(self.device)
        dblobs = []
        with torch.cuda.stream(st):
            for o, b in zip(dev_out, hblobs):
                o[:b.size].copy_(torch.from_numpy(b), non_blocking=True)
                dblobs.append(o[:b.size])
        st.synchronize()
        self.Decompress(dblobs, ltol, s, outs=dev_in)
        with torch.cuda.stream(st):
            for d, a in zip(dev_in, outs):
                torch.from_numpy(a).copy_(d, non_blocking=True)
        st.synchronize()
        return outs

    # ---- one self-describing stream from all ranks -----------------------------------
    def gather_stream(self, blobs, owned, shape, ebtype, tol, s, norm, dd_dim, dd_size,
                      dst=0):
        """header + [u64 size][payload] of every subdomain in subdomain order on rank
        ``dst`` (GPUPipelines.hpp:187-191, Metadata.cpp:353-372): what mgard_x::compress
        returns for the whole domain, decodable by mgx_x_decompress / mgard_x::decompress
        on one GPU.  ``owned`` = global indices of this rank's subdomains."""
        import torch.distributed as dist
        mine = [(int(i), bytes(b.cpu().numpy().tobytes() if hasattr(b, "cpu") else bytes(b)))
                for i, b in zip(owned, blobs)]
        if self.group is not False and dist.is_available() and dist.is_initialized():
            world = dist.get_world_size(self.group)
            rank = dist.get_rank(self.group)
            objs = [None] * world if rank == dst else None
            dist.gather_object(mine, objs, dst=dst, group=self.group)
            if rank != dst:
                return None
            mine = [p for o in objs for p in o]
        mine.sort(key=lambda p: p[0])
        assert [i for i, _ in mine] == list(range(self.num_subdomains)), "missing subdomains"
        head = serialize_header(shape, self.dtype_code, ebtype, tol, s, norm, dd_dim, dd_size)
        return head + frame([p for _, p in mine])
