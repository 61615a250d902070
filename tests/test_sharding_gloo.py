"""N > 1 path on CPU: two gloo ranks run the sharded host logic
(mgard_b200/sharding.py) -- MaxDim decomposition, the single scalar all-reduce,
the local tolerance and the [u64 size][payload] framing -- with the CPU oracle
standing in for the per-rank GPU compressor.  The result must decode, slab by
slab, to a field that meets the GLOBAL error bound."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

HERE = os.path.dirname(os.path.abspath(__file__))
INF = float("inf")


class OracleCompressor:
    """Duck-typed stand-in for mgard_b200.Compressor on one subdomain."""

    def __init__(self, shape, dtype):
        sys.path.insert(0, HERE)
        from oracle_lib import Oracle
        self.o = Oracle("port")
        self.shape, self.dtype = tuple(shape), np.dtype(dtype)

    def NormPartial(self, local, s):
        a = np.asarray(local, np.float64)
        return float(np.abs(a).max()) if np.isinf(s) else float((a * a).sum())

    def Compress(self, local, ebtype, tol, s, norm=0.0, out=None):
        blob, nrm = self.o.compress(np.asarray(local), 1 if ebtype == 0 else 0, tol, s)
        return blob, nrm

    def Decompress(self, blob, ebtype, tol, s, norm, out=None):
        return self.o.decompress(np.asarray(blob), self.shape, self.dtype,
                                 1 if ebtype == 0 else 0, tol, s, norm)


def free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def worker(rank, world, port, shape, dtype, s, ebtype, tol, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        sys.path.insert(0, os.path.dirname(HERE))
        sys.path.insert(0, HERE)
        from oracle_lib import smooth_field
        from mgard_b200.sharding import (ABS, REL, DomainDecomposer,
                                         ShardedCompressor, unframe)
        u = smooth_field(shape, dtype, seed=11)
        u[tuple(0 for _ in shape)] = 7.0  # the global max lives on rank 0's slab only
        dd = DomainDecomposer(shape, world)
        assert dd.num_subdomains == world
        sub_shape, off = dd.subdomain_shape(rank)
        local = np.ascontiguousarray(u[dd.subdomain_slice(rank)])
        assert local.shape == tuple(sub_shape)
        comp = OracleCompressor(sub_shape, dtype)
        sc = ShardedCompressor(comp, 0 if np.dtype(dtype) == np.float32 else 1,
                               u.size, dd.num_subdomains)
        payload, norm, ltol = sc.Compress(local, ebtype, tol, s)
        # every rank sees the same global norm / tolerance
        t = torch.tensor([norm, ltol], dtype=torch.float64)
        lo, hi = t.clone(), t.clone()
        dist.all_reduce(lo, op=dist.ReduceOp.MIN)
        dist.all_reduce(hi, op=dist.ReduceOp.MAX)
        assert torch.equal(lo, hi)
        framed = sc.gather_frames(payload.tobytes(), dst=0)
        back = sc.Decompress(payload, ltol, s)
        err = float(np.abs(back.astype(np.float64) - local).max())
        if rank == 0:
            parts = unframe(framed)
            assert len(parts) == world
            assert parts[0] == payload.tobytes()
            # rank 0 decodes the other ranks' frames too (independent units)
            full = np.empty(shape, dtype)
            for i, p in enumerate(parts):
                shp, _ = dd.subdomain_shape(i)
                c = OracleCompressor(shp, dtype)
                full[dd.subdomain_slice(i)] = c.Decompress(
                    np.frombuffer(p, np.uint8), ABS, ltol, s, 0.0)
            q.put(("ok", norm, ltol, float(np.abs(full.astype(np.float64) - u).max()),
                   float(np.abs(u.astype(np.float64)).max()),
                   float(np.sqrt((u.astype(np.float64) ** 2).sum() / u.size))))
        assert np.isfinite(err)
    except Exception as e:  # pragma: no cover
        if rank == 0:
            q.put(("fail", repr(e)))
        raise
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("shape,dtype,s,ebtype,tol", [
    ((33, 20, 17), np.float32, INF, 0, 1e-3),   # REL L-inf: MAX all-reduce
    ((40, 33), np.float64, 0.0, 0, 1e-3),       # REL L2: SUM all-reduce
    ((33, 20, 17), np.float32, INF, 1, 1e-2),   # ABS: no norm needed
])
def test_two_rank_sharded_compress(shape, dtype, s, ebtype, tol):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = free_port()
    procs = [ctx.Process(target=worker,
                         args=(r, 2, port, shape, dtype, s, ebtype, tol, q))
             for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(180)
        assert p.exitcode == 0
    res = q.get(timeout=10)
    assert res[0] == "ok", res
    _, norm, ltol, err, umax, l2 = res
    if ebtype == 0 and np.isinf(s):
        assert norm == umax                      # max over BOTH slabs
        assert abs(ltol - tol * norm) <= 1e-6 * tol * norm
        assert err <= tol * umax
    elif ebtype == 0:
        assert abs(norm - l2) <= 1e-12 * l2
        assert abs(ltol - np.sqrt((tol * norm) ** 2 / 2)) <= 1e-15
    else:
        assert ltol == pytest.approx(tol, rel=1e-6)
        assert err <= tol


# ---- several subdomains per rank + one self-describing stream ---------------------
def domain_worker(rank, world, port, shape, dtype, s, tol, nsub, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        sys.path.insert(0, os.path.dirname(HERE))
        sys.path.insert(0, HERE)
        from oracle_lib import smooth_field
        from mgard_b200 import _lib
        from mgard_b200.sharding import (ABS, REL, DomainDecomposer, ShardedDomain, unframe)
        from mgard_b200.highlevel import read_metadata
        u = smooth_field(shape, dtype, seed=5)
        dd = DomainDecomposer(shape, nsub)
        S = dd.num_subdomains
        per = -(-S // world)
        owned = [i for i in range(S) if i // per == rank]
        locals_ = [np.ascontiguousarray(u[dd.subdomain_slice(i)]) for i in owned]
        tdt = torch.float32 if np.dtype(dtype) == np.float32 else torch.float64
        dom = ShardedDomain([a.shape for a in locals_], tdt, u.size, S,
                            comp_factory=lambda shp: OracleCompressor(shp, dtype))
        blobs, norm, ltol = dom.Compress(locals_, REL, tol, s)
        stream = dom.gather_stream(blobs, owned, shape, REL, tol, s, norm, dd.dim, dd.size)
        if rank == 0:
            md, hb = read_metadata(stream)
            assert md.decomposed == 1 and md.dd_dim == dd.dim and md.dd_size == dd.size
            as_t = (lambda x: float(np.float32(x))) if np.dtype(dtype) == np.float32 \
                else float
            assert md.norm == as_t(norm) and md.tol == as_t(tol) and md.D == len(shape)
            parts = unframe(stream[hb:])
            assert len(parts) == S
            # the tolerance a decoder derives from the header is the one that was used
            lib = _lib.load()
            assert float(lib.mgx_local_abs_tol(dom.dtype_code, 1, md.norm, md.tol, md.s, S)) == ltol
            full = np.empty(shape, dtype)
            for i, p in enumerate(parts):
                shp, _ = dd.subdomain_shape(i)
                full[dd.subdomain_slice(i)] = OracleCompressor(shp, dtype).Decompress(
                    np.frombuffer(p, np.uint8), ABS, ltol, s, 0.0)
            q.put(("ok", float(np.abs(full.astype(np.float64) - u).max()),
                   float(np.abs(u.astype(np.float64)).max()), S))
    except Exception as e:  # pragma: no cover
        if rank == 0:
            q.put(("fail", repr(e)))
        raise
    finally:
        dist.destroy_process_group()


def test_two_ranks_four_subdomains_one_mgard_stream():
    shape, dtype, tol = (41, 20, 17), np.float32, 1e-3
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = free_port()
    procs = [ctx.Process(target=domain_worker,
                         args=(r, 2, port, shape, dtype, INF, tol, 4, q)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(180)
        assert p.exitcode == 0
    res = q.get(timeout=10)
    assert res[0] == "ok", res
    _, err, umax, S = res
    assert S == 4
    assert err <= tol * umax
