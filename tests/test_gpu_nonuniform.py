"""N2 on the GPU -- non-uniform Cartesian grids (mgx_plan_grid, mgx_x_compress_coords;
Hierarchy(shape, coords, config), Hierarchy.hpp:22-140, 840-857; compress_x.hpp:74-100)
and uniform grids with normalize_coordinates = false, against the oracle (pinned to the
reference built with the same coordinates in tests/test_nonuniform.py)."""
import numpy as np
import pytest
import torch

from oracle_lib import Oracle, smooth_field, grid, same_container
from test_nonuniform import make_coords

import mgard_b200 as mg
from mgard_b200 import highlevel as hl

pytestmark = pytest.mark.gpu
INF = float("inf")


def make(shape, dtype, coords=None, normalize=True):
    tdt = torch.float32 if dtype == np.float32 else torch.float64
    cfg = mg.Config(normalize_coordinates=normalize)
    return mg.Compressor(mg.Hierarchy(shape, tdt, cfg, coords=coords))


SHAPES = [(17,), (18,), (1000,), (9, 12), (65, 130), (5, 6, 9), (33, 40, 65),
          (66, 70, 130), (5, 4, 9, 7), (3, 5, 5, 9, 9)]


@pytest.mark.parametrize("shape", SHAPES)
@pytest.mark.parametrize("dtype", [np.float32, np.float64])
@pytest.mark.parametrize("orth", [0, 1])
def test_decompose_recompose_bit_exact_on_nonuniform_grids(shape, dtype, orth):
    if len(shape) > 3 and not orth:
        pytest.skip("D > 3 always applies the correction")
    u = smooth_field(shape, dtype, seed=3)
    coords = make_coords(shape, 1)
    with grid(coords, dtype=dtype):
        want = Oracle("port").decompose(u, orth)
        back = Oracle("port").recompose(want, orth)
    comp = make(shape, dtype, coords)
    d = torch.from_numpy(u).cuda()
    comp.Decompose(d, orth)
    assert np.array_equal(d.cpu().numpy(), want)
    comp.Recompose(d, orth)
    assert np.array_equal(d.cpu().numpy(), back)


@pytest.mark.parametrize("shape,dtype,s,tol", [
    ((1000,), np.float32, INF, 1e-3), ((18,), np.float64, 0.0, 1e-3),
    ((65, 130), np.float32, INF, 1e-3), ((40, 31), np.float64, 0.0, 1e-4),
    ((33, 40, 65), np.float32, INF, 1e-4), ((66, 70, 130), np.float32, INF, 1e-3),
    ((20, 11, 6), np.float64, 1.0, 1e-3), ((5, 4, 9, 7), np.float64, 0.0, 1e-3)])
def test_containers_equal_the_oracle_on_nonuniform_grids(shape, dtype, s, tol):
    u = smooth_field(shape, dtype, seed=4)
    coords = make_coords(shape, 2)
    with grid(coords, dtype=dtype):
        want, wnorm = Oracle("port").compress(u, 1, tol, s)
    comp = make(shape, dtype, coords)
    d = torch.from_numpy(u).cuda()
    out = torch.empty(4 * u.nbytes + (1 << 20), dtype=torch.uint8, device="cuda")
    blob, norm = comp.Compress(d, mg.REL, tol, s, norm=wnorm, out=out)
    assert same_container(blob.cpu().numpy(), want)
    dec = comp.Decompress(blob, mg.REL, tol, s, wnorm).cpu().numpy().astype(np.float64)
    err = dec - u
    assert (np.abs(err).max() if np.isinf(s) else np.sqrt((err ** 2).mean())) <= tol * wnorm


@pytest.mark.parametrize("shape,dtype,s", [((33,), np.float32, 0.0),
                                           ((17, 20), np.float64, 0.0),
                                           ((33, 40, 65), np.float32, INF),
                                           ((6, 10, 12), np.float64, 1.0)])
def test_unit_spacing_without_normalized_coordinates(shape, dtype, s):
    u = smooth_field(shape, dtype, seed=6)
    with grid(None, normalize=False):
        want, wnorm = Oracle("port").compress(u, 1, 1e-3, s)
    comp = make(shape, dtype, normalize=False)
    d = torch.from_numpy(u).cuda()
    out = torch.empty(4 * u.nbytes + (1 << 20), dtype=torch.uint8, device="cuda")
    blob, norm = comp.Compress(d, mg.REL, 1e-3, s, out=out)  # the GPU's own norm
    if dtype == np.float64 or np.isinf(s):
        assert norm == pytest.approx(wnorm, rel=1e-12)
    blob, _ = comp.Compress(d, mg.REL, 1e-3, s, norm=wnorm, out=out)
    assert same_container(blob.cpu().numpy(), want)


@pytest.mark.parametrize("shape,dtype,s,tol,where", [
    ((33, 40, 65), np.float32, INF, 1e-3, "host"),
    ((33, 40, 65), np.float32, INF, 1e-3, "device"),
    ((40, 50), np.float64, 0.0, 1e-4, "host"),
    ((5, 4, 9, 7), np.float64, 1.0, 1e-2, "host")])
def test_high_level_round_trip_with_coordinates(shape, dtype, s, tol, where):
    u = smooth_field(shape, dtype, seed=8)
    coords = make_coords(shape, 5)
    data = torch.from_numpy(u).cuda() if where == "device" else u
    stream = hl.compress(data, tol, s, hl.REL, coords=coords)
    host = stream.cpu().numpy() if where == "device" else np.asarray(stream)
    md, _ = hl.read_metadata(bytes(host[:65536]))
    assert md.nonuniform == 1
    got = hl.read_coords(host)
    npdt = np.float32 if dtype == np.float32 else np.float64
    for g, c in zip(got, coords):
        assert np.array_equal(g, np.asarray(c, npdt).astype(np.float64))
    back = hl.decompress(stream)
    back = back.cpu().numpy() if where == "device" else back
    err = back.astype(np.float64) - u
    bound = tol * md.norm
    assert (np.abs(err).max() if np.isinf(s) else np.sqrt((err ** 2).mean())) <= bound
    # the stream differs from the uniform-grid one
    uni = hl.compress(data, tol, s, hl.REL)
    uni = uni.cpu().numpy() if where == "device" else np.asarray(uni)
    assert uni.size != host.size or not np.array_equal(uni, host)


def test_high_level_decomposed_with_coordinates():
    """MaxDim subdomains take their slice of the decomposed dimension's coordinates
    (DomainDecomposer.hpp:259-289)"""
    import ctypes
    from mgard_b200 import _lib
    shape, dtype = (70, 20, 24), np.float32
    u = smooth_field(shape, dtype, seed=9)
    coords = make_coords(shape, 6)
    lib = _lib.load()
    shp = (ctypes.c_uint64 * 3)(18, 20, 24)
    fit = lib.mgx_estimate_footprint(3, shp, 0, ctypes.c_double(INF), 1) + 1
    cfg = hl.Config(max_memory_footprint=int(fit))
    stream = hl.compress(u, 1e-3, INF, hl.REL, config=cfg, coords=coords)
    md, _ = hl.read_metadata(bytes(stream[:65536]))
    assert md.decomposed and md.nonuniform
    back = hl.decompress(stream)
    assert np.abs(back.astype(np.float64) - u).max() <= 1e-3 * md.norm


def test_coordinates_must_increase():
    shape = (9, 9)
    c = make_coords(shape, 1)
    c[1][4] = c[1][3]
    with pytest.raises(mg.MgardError):
        make(shape, np.float32, c)
