"""The stream of a domain-decomposed, sharded compress: MGARD header + one
[u64 size][payload] record per MaxDim subdomain (GPUPipelines.hpp:187-191,
Metadata.cpp:353-372), assembled from per-subdomain payloads the way N ranks produce
them, must be byte-identical to what mgx_x_compress (= mgard_x::compress) writes for the
whole domain under the same decomposition, decode on ONE GPU with mgx_x_decompress, and
decode record by record with the reference."""
import ctypes

import numpy as np
import pytest
import torch

import oracle_lib
from oracle_lib import Oracle, smooth_field

import mgard_b200 as mg
from mgard_b200 import _lib, highlevel as hl
from mgard_b200.sharding import ABS, REL, DomainDecomposer, ShardedDomain, unframe

pytestmark = pytest.mark.gpu
INF = float("inf")


def _u64(seq):
    return (ctypes.c_uint64 * len(seq))(*[int(x) for x in seq])


@pytest.mark.parametrize("shape,dtype,s,tol,nsub", [
    ((65, 72, 80), np.float32, INF, 1e-3, 4),      # cfg 5 in small: slabs along axis 0
    ((130, 33, 40), np.float64, 0.0, 1e-3, 8),     # L2: SUM all-reduce, 9-row leftover
    ((6, 8, 9, 9, 200), np.float64, 1.0, 1e-2, 4),  # cfg 4 in small: slabs along the LAST axis
])
def test_sharded_stream_equals_the_single_call_stream(shape, dtype, s, tol, nsub):
    lib = _lib.load()
    u = smooth_field(shape, dtype)
    tdt = torch.float32 if dtype == np.float32 else torch.float64
    code = 0 if dtype == np.float32 else 1
    dd = DomainDecomposer(shape, nsub)
    S = dd.num_subdomains
    assert S >= nsub
    d = torch.from_numpy(u).cuda()
    fields = [d[dd.subdomain_slice(i)].contiguous() for i in range(S)]
    dom = ShardedDomain([tuple(f.shape) for f in fields], tdt, u.size, S,
                        device=torch.device("cuda", 0))
    blobs, norm, ltol = dom.Compress(fields, REL, tol, s)
    stream = dom.gather_stream(blobs, list(range(S)), shape, REL, tol, s, norm, dd.dim, dd.size)
    arr = np.frombuffer(stream, np.uint8).copy()
    # (1) one GPU decodes the N-rank stream through the drop-in entry
    back = hl.decompress(arr)
    assert back.shape == tuple(shape)
    diff = back.astype(np.float64) - u.astype(np.float64)
    if np.isinf(s):
        assert np.abs(diff).max() <= tol * norm
    else:
        assert np.sqrt((diff ** 2).mean()) <= tol * norm
    # (2) the reference (restatement pinned to it) decodes every record on its own
    md, hb = hl.read_metadata(stream)
    parts = unframe(stream[hb:])
    assert len(parts) == S
    o = Oracle("port")
    for i, p in enumerate(parts):
        shp, _ = dd.subdomain_shape(i)
        rec = o.decompress(np.frombuffer(p, np.uint8), shp, dtype, 0, ltol, s, 0.0)
        ref = u[dd.subdomain_slice(i)]
        if np.isinf(s):
            assert np.abs(rec.astype(np.float64) - ref).max() <= tol * norm
    # (3) byte equality with the single-call high-level path under the same decomposition:
    #     a memory budget that one subdomain (with its prefetch buffers,
    #     DomainDecomposer.hpp:293-330) just fits; the halving rule then stops at dd.size
    shp0, _ = dd.subdomain_shape(0)
    fit = lib.mgx_estimate_footprint(len(shape), _u64(shp0), code, ctypes.c_double(s), 1) + 1
    cfg = hl.Config(max_memory_footprint=int(fit))
    single = hl.compress(u, tol, s, hl.REL, config=cfg)
    md1, _ = hl.read_metadata(bytes(single[:4096]))
    assert md1.decomposed and md1.dd_dim == dd.dim and md1.dd_size == dd.size, \
        (md1.decomposed, md1.dd_dim, md1.dd_size, dd.dim, dd.size)
    md0, _ = hl.read_metadata(stream[:4096])
    for f, _t in md0._fields_:
        a, b = getattr(md0, f), getattr(md1, f)
        a = list(a) if hasattr(a, "__len__") else a
        b = list(b) if hasattr(b, "__len__") else b
        assert a == b, "header field %s: sharded %r, single call %r" % (f, a, b)
    assert bytes(single) == stream, "sharded stream differs from mgx_x_compress"
