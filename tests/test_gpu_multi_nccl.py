"""Two NCCL ranks, one GPU each (skipped on a one-GPU box): the ranks compress their
MaxDim subdomains after ONE scalar all-reduce, rank 0 gathers the records behind one
MGARD header (GPUPipelines.hpp:187-191, Metadata.cpp:353-372) and

* the stream is byte-identical to the one a single GPU produces for all subdomains,
* mgx_x_decompress decodes it on rank 0's GPU alone within the GLOBAL bound,
* the oracle (pinned to the reference) decodes every record on its own.
"""
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
pytestmark = pytest.mark.gpu


def free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def worker(rank, world, port, shape, dtype, s, tol, nsub, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    try:
        sys.path.insert(0, os.path.dirname(HERE))
        sys.path.insert(0, HERE)
        from oracle_lib import Oracle, smooth_field
        from mgard_b200 import highlevel as hl
        from mgard_b200.sharding import REL, DomainDecomposer, ShardedDomain, unframe
        u = smooth_field(shape, dtype, seed=5)
        u[tuple(0 for _ in shape)] = 5.0  # the global extreme lives on rank 0 only
        tdt = torch.float32 if dtype == np.float32 else torch.float64
        dd = DomainDecomposer(shape, nsub)
        S = dd.num_subdomains
        per = -(-S // world)
        owned = [i for i in range(S) if i // per == rank]
        fields = [torch.from_numpy(np.ascontiguousarray(u[dd.subdomain_slice(i)])).to(dev)
                  for i in owned]
        dom = ShardedDomain([tuple(f.shape) for f in fields], tdt, u.size, S,
                            device=dev, dev_id=rank)
        blobs, norm, ltol = dom.Compress(fields, REL, tol, s)
        stream = dom.gather_stream(blobs, owned, shape, REL, tol, s, norm, dd.dim, dd.size)
        if rank == 0:
            # all subdomains on one GPU, no process group involved
            allf = [torch.from_numpy(np.ascontiguousarray(u[dd.subdomain_slice(i)])).to(dev)
                    for i in range(S)]
            solo = ShardedDomain([tuple(f.shape) for f in allf], tdt, u.size, S,
                                 group=False, device=dev, dev_id=rank)
            b1, n1, l1 = solo.Compress(allf, REL, tol, s)
            assert n1 == norm and l1 == ltol, (n1, norm, l1, ltol)
            head = len(stream) - sum(8 + int(b.numel()) for b in b1)
            recs = unframe(stream[head:])
            assert [bytes(b.cpu().numpy()) for b in b1] == [bytes(r) for r in recs], \
                "two-rank stream differs from the one-GPU stream"
            back = hl.decompress(np.frombuffer(stream, np.uint8).copy(),
                                 config=hl.Config(dev_id=0))
            diff = back.astype(np.float64) - u.astype(np.float64)
            err = np.abs(diff).max() if np.isinf(s) else np.sqrt((diff ** 2).mean())
            assert err <= tol * norm, (err, tol * norm)
            o = Oracle("port")
            for i, r in enumerate(recs):
                shp, _ = dd.subdomain_shape(i)
                rec = o.decompress(np.frombuffer(r, np.uint8), shp, dtype, 0, ltol, s, 0.0)
                ref = u[dd.subdomain_slice(i)].astype(np.float64)
                d = rec.astype(np.float64) - ref
                if np.isinf(s):
                    assert np.abs(d).max() <= tol * norm
        q.put((rank, "ok"))
    except Exception as e:  # noqa: BLE001
        import traceback
        q.put((rank, "FAIL: " + repr(e) + "\n" + traceback.format_exc()))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("shape,dtype,s,tol,nsub", [
    ((130, 72, 80), np.float32, INF, 1e-3, 4),     # cfg 5 in small: slabs along axis 0
    ((6, 8, 9, 9, 200), np.float64, 1.0, 1e-2, 4),  # cfg 4 in small: slabs along the last axis
])
def test_two_nccl_ranks_one_mgard_stream(shape, dtype, s, tol, nsub):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = free_port()
    procs = [ctx.Process(target=worker, args=(r, 2, port, shape, dtype, s, tol, nsub, q))
             for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=600) for _ in procs]
    for p in procs:
        p.join(60)
    assert all(m == "ok" for _, m in res), res
