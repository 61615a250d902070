"""N4 (Block): Config::domain_decomposition = Block cuts the domain into cubes of
Config::block_size (DomainDecomposer.hpp:95-160, 226-251); every block is its own
compress unit with the ABS bound of calc_local_abs_tol, records in block order (last
dimension fastest) behind a header that says (BLOCK, size).  Checked: the header, the
bound on every element through mgx_x_decompress, every record decoded on its own by the
oracle against the block it came from, host and device buffers, leftover blocks."""
import numpy as np
import pytest
import torch

from oracle_lib import Oracle, smooth_field, grid
from test_nonuniform import make_coords

from mgard_b200 import highlevel as hl
from mgard_b200.sharding import unframe

pytestmark = pytest.mark.gpu
INF = float("inf")
BLOCK = 1  # mgx_dd_type / domain_decomposition_type::Block


def blocks_of(shape, size):
    """[(slices, shape)] in the reference's subdomain order"""
    nb = [(n - 1) // size + 1 for n in shape]
    out = []
    for bid in range(int(np.prod(nb))):
        rem, idx = bid, [0] * len(shape)
        for d in range(len(shape) - 1, -1, -1):
            idx[d] = rem % nb[d]
            rem //= nb[d]
        sl, shp = [], []
        for d, n in enumerate(shape):
            ext = size if idx[d] < n // size else n % size
            sl.append(slice(idx[d] * size, idx[d] * size + ext))
            shp.append(ext)
        out.append((tuple(sl), tuple(shp)))
    return out


@pytest.mark.parametrize("shape,dtype,s,tol,size,where", [
    ((70, 40, 50), np.float32, INF, 1e-3, 32, "host"),     # leftovers 6, 8, 18
    ((70, 40, 50), np.float32, INF, 1e-3, 32, "device"),
    ((64, 96), np.float64, 0.0, 1e-4, 32, "host"),          # exact multiples, L2
    ((20, 24, 11, 12), np.float64, 1.0, 1e-2, 8, "host"),   # 4-D, s-norm
    ((300,), np.float32, INF, 1e-4, 128, "host"),
])
def test_block_decomposition_round_trip(shape, dtype, s, tol, size, where):
    u = smooth_field(shape, dtype, seed=7)
    cfg = hl.Config(domain_decomposition=BLOCK, block_size=size)
    data = torch.from_numpy(u).cuda() if where == "device" else u
    stream = hl.compress(data, tol, s, hl.REL, config=cfg)
    host = stream.cpu().numpy() if where == "device" else np.asarray(stream)
    md, hb = hl.read_metadata(bytes(host[:65536]))
    assert md.decomposed and md.dd_type == BLOCK and md.dd_size == size
    back = hl.decompress(stream)
    back = back.cpu().numpy() if where == "device" else back
    err = back.astype(np.float64) - u
    bound = tol * md.norm
    assert (np.abs(err).max() if np.isinf(s) else np.sqrt((err ** 2).mean())) <= bound
    # every record on its own, by the oracle, against the block it came from
    blocks = blocks_of(shape, size)
    recs = unframe(bytes(host[hb:]))
    assert len(recs) == len(blocks)
    ltol = float(np.float32(tol * md.norm)) if np.isinf(s) else \
        float(np.sqrt((tol * md.norm) ** 2 / len(blocks)))
    o = Oracle("port")
    worst, sq = 0.0, 0.0
    for (sl, shp), rec in zip(blocks, recs):
        ref = u[sl].astype(np.float64)
        r = np.frombuffer(rec, np.uint8)
        if r.size == ref.size * u.itemsize:  # raw fallback
            dec = r.view(u.dtype).reshape(shp).astype(np.float64)
        else:
            dec = o.decompress(r, shp, dtype, 0, ltol, s, 0.0).astype(np.float64)
        d = dec - ref
        worst = max(worst, float(np.abs(d).max()))
        sq += float((d * d).sum())
    assert (worst if np.isinf(s) else np.sqrt(sq / u.size)) <= bound


def test_block_decomposition_with_coordinates():
    shape, size = (40, 36), 16
    u = smooth_field(shape, np.float64, seed=3)
    coords = make_coords(shape, 4)
    cfg = hl.Config(domain_decomposition=BLOCK, block_size=size)
    stream = hl.compress(u, 1e-4, 0.0, hl.REL, config=cfg, coords=coords)
    md, _ = hl.read_metadata(bytes(stream[:65536]))
    assert md.decomposed and md.dd_type == BLOCK and md.nonuniform
    back = hl.decompress(stream)
    assert np.sqrt(((back - u) ** 2).mean()) <= 1e-4 * md.norm


def test_a_block_of_fewer_than_three_nodes_is_refused():
    u = smooth_field((34, 40), np.float32)  # 34 = 32 + 2
    cfg = hl.Config(domain_decomposition=BLOCK, block_size=32)
    with pytest.raises(Exception):
        hl.compress(u, 1e-3, INF, hl.REL, config=cfg)
