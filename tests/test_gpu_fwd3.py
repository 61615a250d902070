"""GPU parity of the TMA-staged 3-D forward pass (mgx_hier3.cu, k_fwd3 + k_fwd3_rest +
k_hist_fix) through the C ABI:

* quantised codes of every level == oracle Decompose(hierarchical) + Quantize,
* whole containers (symbols, histogram -> codebook, outlier list) byte-identical to
  the oracle's and to the LDGSTS kernel family (MGX_NO_TMA=1),
on 2^k+1, even, ragged and slab-like shapes, fp32 and fp64, with bounds that give
all-zero blocks, mixed blocks and outliers.
"""
import numpy as np
import pytest
import torch

import oracle_lib
from oracle_lib import Oracle, smooth_field

import mgard_b200 as mg

pytestmark = pytest.mark.gpu
INF = float("inf")

# (shape, dtype): nz, ny >= 9 and nx >= 65 take the TMA kernel
SHAPES = [((65, 65, 65), np.float32), ((129, 129, 129), np.float32),
          ((65, 65, 65), np.float64), ((73, 81, 200), np.float32),
          ((17, 41, 129), np.float32), ((66, 70, 130), np.float64),
          ((33, 257, 193), np.float32), ((25, 24, 321), np.float64),
          ((9, 9, 65), np.float32), ((130, 20, 67), np.float32)]


def make(shape, dtype):
    tdt = torch.float32 if dtype == np.float32 else torch.float64
    return mg.Compressor(mg.Hierarchy(shape, tdt))


def rough_field(shape, dtype, seed=7):
    """smooth field + strong small-scale noise: mixed and non-zero blocks at every level"""
    rng = np.random.default_rng(seed)
    return (smooth_field(shape, np.float64) + 0.02 * rng.standard_normal(shape)).astype(dtype)


@pytest.mark.parametrize("shape,dtype", SHAPES)
@pytest.mark.parametrize("rel,tol", [(1, 1e-3), (0, 1e-4), (1, 3e-6)])
@pytest.mark.parametrize("kind", ["smooth", "rough"])
def test_codes_equal_the_oracle(shape, dtype, rel, tol, kind):
    u = smooth_field(shape, dtype) if kind == "smooth" else rough_field(shape, dtype)
    o = Oracle("port")
    norm = float(np.abs(u).max())
    want = o.quantize(o.decompose(u, 0), rel, tol, INF, norm)
    comp = make(shape, dtype)
    got = comp.DecomposeQuantizeFused(torch.from_numpy(u).cuda(),
                                      mg.REL if rel else mg.ABS, tol, norm).cpu().numpy()
    bad = np.argwhere(got != want)
    assert len(bad) == 0, "%d codes differ, first %s: got %s want %s" % (
        len(bad), bad[:5].tolist(), got[tuple(bad[0])], want[tuple(bad[0])])


@pytest.mark.parametrize("shape,dtype", SHAPES)
@pytest.mark.parametrize("kind,tol", [("smooth", 1e-3), ("rough", 1e-4), ("rough", 2e-5)])
def test_containers_equal_the_oracle_and_the_ldgsts_kernels(shape, dtype, kind, tol,
                                                            monkeypatch):
    u = smooth_field(shape, dtype) if kind == "smooth" else rough_field(shape, dtype)
    want, wnorm = Oracle("port").compress(u, 1, tol, INF)
    d = torch.from_numpy(u).cuda()
    comp = make(shape, dtype)
    big = lambda: torch.empty(6 * u.nbytes + (1 << 20), dtype=torch.uint8, device="cuda")
    blob, norm = comp.Compress(d, mg.REL, tol, INF, out=big())
    got = blob.cpu().numpy()
    assert norm == wnorm
    assert got.size == want.size and np.array_equal(np.delete(got, 7), np.delete(want, 7)), \
        "container differs from the oracle's"
    monkeypatch.setenv("MGX_NO_TMA", "1")
    comp_old = make(shape, dtype)
    monkeypatch.delenv("MGX_NO_TMA")
    blob_old, _ = comp_old.Compress(d, mg.REL, tol, INF, out=big())
    assert torch.equal(blob, blob_old), "TMA and LDGSTS kernel families disagree"
    # twice in a row on one context (histogram / outlier counters are reset per call)
    blob2, _ = comp.Compress(d, mg.REL, tol, INF, out=big())
    assert torch.equal(blob, blob2)
    dec = comp.Decompress(blob, mg.REL, tol, INF, norm).cpu().numpy()
    assert np.abs(dec.astype(np.float64) - u.astype(np.float64)).max() <= tol * norm


def test_unaligned_input_pointer_takes_the_ldgsts_kernel():
    """a field that starts 4 bytes into an allocation cannot be described by a tensor
    map (16-byte base): same bytes through the fallback"""
    shape = (33, 65, 129)
    u = smooth_field(shape, np.float32)
    want, _ = Oracle("port").compress(u, 1, 1e-3, INF)
    buf = torch.zeros(u.size + 1, dtype=torch.float32, device="cuda")
    view = buf[1:].view(shape)
    view.copy_(torch.from_numpy(u))
    assert view.data_ptr() % 16 == 4
    comp = make(shape, np.float32)
    blob, _ = comp.Compress(view, mg.REL, 1e-3, INF)
    got = blob.cpu().numpy()
    assert got.size == want.size and np.array_equal(np.delete(got, 7), np.delete(want, 7))


# ---- tile draws of several tiles, staged outliers ------------------------------------
def spiky_field(shape, dtype, seed=11):
    """smooth field with narrow peaks of 1e3 x its amplitude: the coefficients around a
    peak leave the dictionary, hundreds of outliers in a few tiles (more than the 32
    slots a warp stages) next to tiles with none"""
    rng = np.random.default_rng(seed)
    u = smooth_field(shape, np.float64)
    for _ in range(24):
        c = [int(rng.integers(4, n - 4)) for n in shape]
        sl = tuple(slice(ci - 3, ci + 4) for ci in c)
        u[sl] += 1e3 * rng.standard_normal(u[sl].shape)
    return u.astype(dtype)


@pytest.mark.parametrize("shape,dtype,draw", [
    ((210, 210, 530), np.float32, 4),    # 26 x 26 x 8 regular tiles: draws of 4
    ((137, 330, 1100), np.float32, 8),   # 17 x 41 x 17 tiles, ragged super blocks: draws of 8
    ((130, 265, 330), np.float64, 2),    # fp64: 16 x 33 x 5 tiles, draws of 2
])
def test_drawn_tiles_and_outlier_staging_match_the_ldgsts_kernels(shape, dtype, draw,
                                                                  monkeypatch):
    """k_fwd3 with tiles drawn several at a time from the device-wide counter and with
    outlier-dense tiles: container bytes equal to the LDGSTS kernel family (one tile
    list per CTA, one global atomic per outlier), which the small cases pin to the
    oracle; twice in a row on one context (tile counter and outlier counter reset)."""
    u = spiky_field(shape, dtype)
    d = torch.from_numpy(u).cuda()
    tol = 1e-6
    big = lambda: torch.empty(3 * u.nbytes + (1 << 20), dtype=torch.uint8, device="cuda")
    comp = make(shape, dtype)
    blob, norm = comp.Compress(d, mg.REL, tol, INF, out=big())
    blob = blob.clone()
    monkeypatch.setenv("MGX_NO_TMA", "1")
    comp_old = make(shape, dtype)
    monkeypatch.delenv("MGX_NO_TMA")
    blob_old, norm_old = comp_old.Compress(d, mg.REL, tol, INF, out=big())
    assert norm == norm_old
    assert torch.equal(blob, blob_old), "TMA and LDGSTS kernel families disagree"
    blob2, _ = comp.Compress(d, mg.REL, tol, INF, out=big())
    assert torch.equal(blob, blob2)
    # the outlier list is not empty, and the stream decodes within the bound
    n = u.size
    nchunk = (n - 1) // 1024 + 1
    host = blob.cpu().numpy()
    off_ddata = 32 + 16 * nchunk + 8 + 8 * (128 + 8192)
    ddata = int(host[off_ddata:off_ddata + 8].view(np.uint64)[0])
    n_out = int(host[off_ddata + 8 + 8 * ddata:off_ddata + 16 + 8 * ddata].view(np.uint64)[0])
    assert n_out > 1000, n_out
    dec = comp.Decompress(blob, mg.REL, tol, INF, norm).cpu().numpy()
    assert np.abs(dec.astype(np.float64) - u.astype(np.float64)).max() <= tol * norm
