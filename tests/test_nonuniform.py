"""N2 -- non-uniform Cartesian grids and normalize_coordinates = false, CPU side:

* the C restatement against the reference itself built with
  Hierarchy(shape, coords, config) (Hierarchy.hpp:22-140, 840-857): decompose /
  recompose bit for bit, whole containers byte for byte;
* uniform grids with unit spacing (Config::normalize_coordinates = false:
  create_uniform_coords, Hierarchy.hpp:762-782; L2 norm without the 1/N,
  NormCalculator.hpp:64);
* node coordinates through the "MGARD" header (Metadata.hpp:138-153,
  Metadata.cpp:303-321, 557-580) -- host-only entry points of the C ABI.
"""
import ctypes

import numpy as np
import pytest

import oracle_lib
from oracle_lib import Oracle, smooth_field, grid, same_container

INF = float("inf")
needs_ref = pytest.mark.skipif(not oracle_lib.have_ref(),
                               reason="oracle/_ref not built (needs /root/reference)")


def make_coords(shape, seed=0):
    """strictly increasing, visibly non-uniform node coordinates in [0, 1]"""
    rng = np.random.default_rng(seed)
    out = []
    for n in shape:
        h = rng.uniform(0.3, 1.7, n - 1)
        c = np.concatenate([[0.0], np.cumsum(h)])
        out.append(c / c[-1])
    return out


SHAPES = [(17,), (18,), (64,), (9, 12), (17, 33), (5, 6, 9), (9, 9, 17), (6, 10, 12),
          (5, 4, 9, 7)]


@needs_ref
@pytest.mark.parametrize("shape", SHAPES)
@pytest.mark.parametrize("dtype", [np.float32, np.float64])
@pytest.mark.parametrize("orth", [0, 1])
def test_restatement_equals_the_reference_on_nonuniform_grids(shape, dtype, orth):
    if len(shape) > 3 and not orth:
        pytest.skip("D > 3 always runs with the orthogonal correction")
    u = smooth_field(shape, dtype, seed=3)
    with grid(make_coords(shape, 1), dtype=dtype):
        r, p = Oracle("ref"), Oracle("port")
        cr, cp = r.decompose(u, orth), p.decompose(u, orth)
        assert np.array_equal(cr, cp)
        assert np.array_equal(r.recompose(cr, orth), p.recompose(cp, orth))
    # and the grid really changed the result
    assert not np.array_equal(cp, Oracle("port").decompose(u, orth))


@needs_ref
@pytest.mark.parametrize("shape,dtype,s,tol", [
    ((64,), np.float32, INF, 1e-3), ((18,), np.float64, 0.0, 1e-3),
    ((17, 33), np.float32, INF, 1e-3), ((9, 12), np.float64, 0.0, 1e-4),
    ((9, 9, 17), np.float32, INF, 1e-4), ((6, 10, 12), np.float64, 1.0, 1e-3),
    ((5, 4, 9, 7), np.float64, 0.0, 1e-3)])
def test_containers_equal_the_reference_on_nonuniform_grids(shape, dtype, s, tol):
    u = smooth_field(shape, dtype, seed=4)
    with grid(make_coords(shape, 2), dtype=dtype):
        want, wnorm = oracle_lib.fresh_in_grid("ref", make_coords(shape, 2), dtype, True,
                                               "compress", u, 1, tol, s)
        got, gnorm = Oracle("port").compress(u, 1, tol, s)
        assert gnorm == wnorm
        assert same_container(got, want)
        back = Oracle("port").decompress(got, shape, dtype, 1, tol, s, gnorm)
    err = back.astype(np.float64) - u
    bound = tol * gnorm
    assert (np.abs(err).max() if np.isinf(s) else np.sqrt((err ** 2).mean())) <= bound


@needs_ref
@pytest.mark.parametrize("shape,dtype,s", [((33,), np.float32, 0.0),
                                           ((17, 20), np.float64, 0.0),
                                           ((9, 9, 9), np.float32, INF),
                                           ((6, 10, 12), np.float64, 1.0)])
def test_unit_spacing_without_normalized_coordinates(shape, dtype, s):
    u = smooth_field(shape, dtype, seed=6)
    with grid(None, normalize=False):
        want, wnorm = oracle_lib.fresh_in_grid("ref", None, dtype, False, "compress", u,
                                               1, 1e-3, s)
        got, gnorm = Oracle("port").compress(u, 1, 1e-3, s)
        assert gnorm == wnorm
        assert same_container(got, want)
    if not np.isinf(s):  # sqrt(sum u^2), not sqrt(sum / N)
        _, nnorm = Oracle("port").compress(u, 1, 1e-3, s)
        assert gnorm == pytest.approx(nnorm * np.sqrt(u.size), rel=1e-5)


# ---- header ------------------------------------------------------------------
def test_coordinates_travel_in_the_header():
    from mgard_b200 import _lib, highlevel as hl
    lib = _lib.load()
    shape = (5, 7, 9)
    coords = make_coords(shape, 3)
    flat = np.concatenate(coords).astype(np.float64)
    md = _lib.MgxMetadata()
    md.D = 3
    for d in range(3):
        md.shape[d] = shape[d]
    md.version[0], md.file_version[0] = 1, 1
    md.nonuniform = 1
    md.dtype, md.mode, md.linf = 1, 1, 0
    md.s, md.tol, md.norm = 0.0, 1e-3, 2.5
    md.decomposition, md.lossless, md.backend = 0, 0, 3
    md.huff_dict_size, md.huff_block_size = 8192, 1024
    out = np.zeros(4096, np.uint8)
    nb = ctypes.c_uint64()
    rc = lib.mgx_metadata_serialize_coords(ctypes.byref(md), flat.ctypes.data,
                                           out.ctypes.data, out.size, ctypes.byref(nb))
    assert rc == 0
    head = out[:nb.value]
    md2, hb = hl.read_metadata(head)
    assert hb == nb.value and md2.nonuniform == 1 and md2.D == 3
    got = hl.read_coords(head)
    assert len(got) == 3 and all(np.array_equal(g, c) for g, c in zip(got, coords))
    # without the coordinates a non-uniform header is refused
    assert lib.mgx_metadata_serialize_coords(ctypes.byref(md), None, out.ctypes.data,
                                             out.size, ctypes.byref(nb)) != 0
    # a uniform header carries none
    md.nonuniform = 0
    assert lib.mgx_metadata_serialize(ctypes.byref(md), out.ctypes.data, out.size,
                                      ctypes.byref(nb)) == 0
    assert hl.read_coords(out[:nb.value]) is None
    # truncated coordinate list: count mismatch
    md.nonuniform = 1
    rc = lib.mgx_metadata_serialize_coords(ctypes.byref(md), flat.ctypes.data,
                                           out.ctypes.data, out.size, ctypes.byref(nb))
    assert rc == 0
