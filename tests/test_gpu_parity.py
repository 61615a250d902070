"""GPU parity tests: every stage of the CUDA path, called through the C ABI,
against the CPU oracle on the same seeded inputs.

Oracle = oracle/libmgx_oracle.so (C restatement, itself pinned bit-for-bit to the
reference in test_oracle_vs_ref.py / test_golden.py) and, when the prebuilt
oracle/_ref/libmgardx_ref.so travelled with the repo, the reference itself.

Bars: coefficients within 1e-5 (fp32) / 1e-12 (fp64) of ||coeff||_inf (they are
in fact bit-identical); quantised codes identical; Huffman container bytes
identical; error bound on every element.
"""
import numpy as np
import pytest
import torch

import oracle_lib
from oracle_lib import Oracle, smooth_field

import mgard_b200 as mg

pytestmark = pytest.mark.gpu
INF = float("inf")


def oracles():
    out = [Oracle("port")]
    if oracle_lib.have_ref():
        out.append(Oracle("ref"))
    return out


def make(shape, dtype):
    tdt = torch.float32 if dtype == np.float32 else torch.float64
    h = mg.Hierarchy(shape, tdt)
    return mg.Compressor(h), tdt


SHAPES = [((33,), np.float32), ((4097,), np.float64), ((40,), np.float32),
          ((17, 17), np.float64), ((129, 129), np.float32),
          ((20, 11), np.float64), ((9, 9, 9), np.float32),
          ((33, 33, 33), np.float64), ((20, 11, 6), np.float32),
          ((65, 70, 130), np.float32), ((5, 6, 7, 8), np.float64),
          ((5, 5, 5, 5, 9), np.float64), ((4, 6, 5, 5, 11), np.float32)]


@pytest.mark.parametrize("shape,dtype", SHAPES)
@pytest.mark.parametrize("orth", [0, 1])
def test_decompose_recompose(shape, dtype, orth):
    if len(shape) > 3 and not orth:
        pytest.skip("D > 3 always applies the correction")
    u = smooth_field(shape, dtype)
    comp, tdt = make(shape, dtype)
    d = torch.from_numpy(u).cuda()
    comp.Decompose(d, orth)
    got = d.cpu().numpy()
    tol = 1e-5 if dtype == np.float32 else 1e-12
    for o in oracles():
        want = o.decompose(u, orth)
        scale = np.abs(want).max()
        assert np.abs(got - want).max() <= tol * scale
        # stronger, observed: bit-identical
        assert np.array_equal(got, want), "coefficients differ from " + o.kind
    comp.Recompose(d, orth)
    back = d.cpu().numpy()
    want_back = oracles()[0].recompose(got, orth)
    assert np.array_equal(back, want_back)
    assert np.abs(back - u).max() <= (1e-5 if dtype == np.float32 else 1e-12) * np.abs(u).max()


@pytest.mark.parametrize("shape,dtype", SHAPES)
@pytest.mark.parametrize("s", [INF, 0.0, 1.0])
@pytest.mark.parametrize("rel", [0, 1])
def test_quantize_dequantize(shape, dtype, s, rel):
    u = smooth_field(shape, dtype)
    orth = mg.infer_orthogonal_projection(len(shape), s)
    o = oracles()[-1]
    c = o.decompose(u, orth)
    norm = 0.73
    want = o.quantize(c, rel, 1e-3, s, norm)
    comp, tdt = make(shape, dtype)
    eb = mg.REL if rel else mg.ABS
    got = comp.Quantize(torch.from_numpy(c), eb, 1e-3, s, norm).cpu().numpy()
    mism = int((got != want).sum())
    assert mism == 0, "%d quantised codes differ" % mism
    deq = comp.Dequantize(torch.from_numpy(want), eb, 1e-3, s, norm).cpu().numpy()
    assert np.array_equal(deq, o.dequantize(want, dtype, rel, 1e-3, s, norm))


def _code_cases():
    rng = np.random.default_rng(7)
    cases = []
    for p in (0.02, 0.17, 0.47, 0.9):
        n = 300000
        cases.append(("geom%.2f" % p, rng.geometric(p, n) * rng.choice([-1, 1], n) + 4096))
    cases.append(("uniform", rng.integers(0, 8192, 200000)))
    cases.append(("outliers", rng.integers(-100, 9000, 50001)))
    cases.append(("single", np.full(5000, 4096)))
    cases.append(("normal30", rng.normal(0, 30, 1234567).astype(np.int64) + 4096))
    cases.append(("normal300", rng.normal(0, 300, 1000003).astype(np.int64) + 4096))
    cases.append(("tiny", np.array([4096, 4097, 4096])))
    return cases


@pytest.mark.parametrize("name,codes", _code_cases(), ids=[c[0] for c in _code_cases()])
def test_huffman_container_bit_exact(name, codes):
    codes = np.ascontiguousarray(codes, np.int64)
    n = codes.size
    comp, _ = make((n,), np.float32) if n >= 3 else (None, None)
    o = Oracle("port")
    want = o.huffman_compress(codes)
    got = comp.LosslessCompress(torch.from_numpy(codes)).cpu().numpy()
    assert got.size == want.size
    assert np.array_equal(got, want), "container bytes differ (first at %d)" % int(
        np.nonzero(got != want)[0][0])
    # our decoder on the oracle's stream, the oracle's decoder on ours
    back = comp.LosslessDecompress(torch.from_numpy(want)).cpu().numpy()
    assert np.array_equal(back, codes)
    assert np.array_equal(o.huffman_decompress(got, n), codes)


@pytest.mark.parametrize("seed", range(12))
def test_codebook_random_histograms(seed):
    rng = np.random.default_rng(100 + seed)
    k = int(rng.integers(2, 8192))
    n = int(rng.integers(k, 400000))
    sc = rng.uniform(0.5, k / 3 + 1)
    codes = np.clip(rng.laplace(0, sc, n).astype(np.int64) + k // 2, 0, k - 1)
    comp, _ = make((n,), np.float32)
    want = Oracle("port").huffman_compress(codes)
    got = comp.LosslessCompress(torch.from_numpy(codes)).cpu().numpy()
    assert np.array_equal(got, want)


E2E = [((65, 65, 65), np.float32, INF, 1, 1e-3), ((129, 129), np.float64, 0.0, 1, 1e-4),
       ((4097,), np.float32, INF, 0, 1e-4), ((8, 16, 9, 9, 33), np.float64, 1.0, 1, 1e-3),
       ((40, 30, 21), np.float32, INF, 1, 1e-2), ((40, 30, 21), np.float64, 0.0, 0, 1e-3),
       ((33, 33, 33), np.float32, INF, 1, 1e-6), ((129, 129, 129), np.float32, INF, 1, 1e-3),
       ((100, 257, 70), np.float64, INF, 1, 1e-5), ((513, 513), np.float32, INF, 1, 1e-4)]


@pytest.mark.parametrize("shape,dtype,s,rel,tol", E2E)
def test_compress_decompress_vs_oracle(shape, dtype, s, rel, tol):
    u = smooth_field(shape, dtype)
    o = Oracle("port")
    want, norm = o.compress(u, rel, tol, s)
    if oracle_lib.have_ref():
        # the reference itself, from a fresh process (see oracle_lib.fresh)
        rwant, rnorm = oracle_lib.fresh("ref", "compress", u, rel, tol, s)
        allowed = oracle_lib.ref_allowed_outputs("compress", u, rel, tol, s)
        assert rnorm == norm
        assert any(oracle_lib.same_container(rwant, a[0]) for a in allowed)
    comp, tdt = make(shape, dtype)
    eb = mg.REL if rel else mg.ABS
    # L2 norms accumulate in fp64 on the GPU (T on the SERIAL reference): hand
    # the oracle's norm over so that byte parity is about the codec only
    blob, norm_gpu = comp.Compress(torch.from_numpy(u).cuda(), eb, tol, s,
                                   norm=norm if rel else 0.0)
    got = blob.cpu().numpy()
    if rel:
        own = comp.CalculateNorm(torch.from_numpy(u).cuda(), s)
        assert abs(own - norm) <= 1e-6 * abs(norm)
    assert got.size == want.size, "compressed size %d vs oracle %d" % (got.size, want.size)
    assert np.array_equal(np.delete(got, 7), np.delete(want, 7))
    dec = comp.Decompress(blob, eb, tol, s, norm).cpu().numpy()
    ref_dec = o.decompress(want, shape, dtype, rel, tol, s, norm)
    assert np.array_equal(dec, ref_dec)
    if np.isinf(s):
        bound = tol * (np.abs(u).max() if rel else 1.0)
        assert np.abs(dec.astype(np.float64) - u).max() <= bound


def test_host_pointer_round_trip():
    shape, dtype = (65, 65, 65), np.float32
    u = smooth_field(shape, dtype)
    comp, _ = make(shape, dtype)
    blob, norm = comp.CompressHost(u, mg.REL, 1e-3, INF)
    want, wnorm = Oracle("port").compress(u, 1, 1e-3, INF)
    assert norm == wnorm
    assert np.array_equal(np.delete(blob, 7), np.delete(want, 7))
    dec = comp.DecompressHost(blob, mg.REL, 1e-3, INF, norm)
    assert np.abs(dec - u).max() <= 1e-3 * np.abs(u).max()


def test_error_paths():
    with pytest.raises(mg.MgardError):
        mg.Compressor(mg.Hierarchy((2, 9), torch.float32))  # dim < 3
    with pytest.raises(mg.MgardError):
        mg.Hierarchy((3,) * 6, torch.float32)  # D > 5
    comp, _ = make((33, 33), np.float32)
    u = torch.from_numpy(smooth_field((33, 33), np.float32)).cuda()
    small = torch.empty(64, dtype=torch.uint8, device="cuda")
    with pytest.raises(mg.MgardError) as e:
        comp.Compress(u, mg.REL, 1e-3, INF, out=small)
    assert e.value.status == 2  # OutputTooLargeFailure
    bad = torch.zeros(4096, dtype=torch.uint8, device="cuda")
    with pytest.raises(mg.MgardError):
        comp.Decompress(bad, mg.REL, 1e-3, INF, 1.0)  # signature mismatch


FUSED = [((33,), np.float32), ((4097,), np.float64), ((40,), np.float32),
         ((17, 17), np.float64), ((129, 129), np.float32), ((20, 11), np.float64),
         ((513, 513), np.float32), ((9, 9, 9), np.float32), ((33, 33, 33), np.float64),
         ((20, 11, 6), np.float32), ((65, 65, 65), np.float32), ((65, 70, 130), np.float32),
         ((100, 257, 70), np.float64)]


@pytest.mark.parametrize("shape,dtype", FUSED)
@pytest.mark.parametrize("rel,tol", [(1, 1e-3), (0, 1e-4), (1, 1e-6)])
def test_fused_hierarchical_forward(shape, dtype, rel, tol):
    """one-pass coefficients+quantise == oracle Decompose(hier) + Quantize"""
    u = smooth_field(shape, dtype)
    o = oracles()[-1]
    norm = float(np.abs(u).max())
    want = o.quantize(o.decompose(u, 0), rel, tol, INF, norm)
    comp, _ = make(shape, dtype)
    got = comp.DecomposeQuantizeFused(torch.from_numpy(u).cuda(),
                                      mg.REL if rel else mg.ABS, tol, norm).cpu().numpy()
    bad = np.argwhere(got != want)
    assert len(bad) == 0, "%d codes differ, first %s: got %s want %s" % (
        len(bad), bad[:5].tolist(), got[tuple(bad[0])], want[tuple(bad[0])])


# ---- row-streaming level kernels (mgx_level.cu): orthogonal-projection path ----
ORTH_SHAPES = [((1025,), np.float64), ((300, 77), np.float64), ((257, 513), np.float32),
               ((66, 33, 130), np.float64), ((31, 64, 65), np.float32),
               ((6, 7, 9, 70), np.float32), ((5, 6, 7, 8, 40), np.float64)]


@pytest.mark.parametrize("shape,dtype", ORTH_SHAPES)
def test_level_kernels_vs_oracle_and_gather_kernels(shape, dtype, monkeypatch):
    """decompose / recompose with the correction: the row-streaming kernels
    against the oracle AND against the superseded gather kernels
    (MGX_LEGACY_LEVELS=1), bit for bit; even sizes, ragged rows, 1-5 D."""
    rng = np.random.default_rng(7)
    u = (smooth_field(shape, dtype) + 0.01 * rng.standard_normal(shape)).astype(dtype)
    comp, tdt = make(shape, dtype)
    d = torch.from_numpy(u).cuda()
    comp.Decompose(d, 1)
    got = d.cpu().numpy()
    want = Oracle("port").decompose(u, 1)
    assert np.array_equal(got, want)
    monkeypatch.setenv("MGX_LEGACY_LEVELS", "1")
    old, _ = make(shape, dtype)
    monkeypatch.delenv("MGX_LEGACY_LEVELS")
    d2 = torch.from_numpy(u).cuda()
    old.Decompose(d2, 1)
    assert np.array_equal(d2.cpu().numpy(), got)
    comp.Recompose(d, 1)
    old.Recompose(d2, 1)
    assert np.array_equal(d.cpu().numpy(), d2.cpu().numpy())
    assert np.array_equal(d.cpu().numpy(), Oracle("port").recompose(want, 1))


@pytest.mark.parametrize("scale", [1e-300, 1e-280, 1e280, 0.0])
def test_thomas_fast_division_falls_back_outside_its_range(scale):
    """The Thomas solves divide through a precomputed reciprocal with two
    exact-residual corrections; dividends outside 2^-897..2^897 (and the
    batches they sit in) must take the true division.  Mixed magnitudes and
    exact zeros against the oracle, bit for bit."""
    shape, dtype = (129, 65), np.float64
    u = smooth_field(shape, dtype)
    u[::3, :] *= scale if scale else 0.0
    u[5:40, 7:20] = 0.0
    comp, tdt = make(shape, dtype)
    d = torch.from_numpy(u).cuda()
    comp.Decompose(d, 1)
    want = Oracle("port").decompose(u, 1)
    assert np.array_equal(d.cpu().numpy(), want)
    comp.Recompose(d, 1)
    assert np.array_equal(d.cpu().numpy(), Oracle("port").recompose(want, 1))


ORTH_E2E = [((300, 77), np.float64, 0.0, 1, 1e-4), ((257, 513), np.float32, 1.0, 1, 1e-3),
            ((66, 33, 130), np.float32, 0.0, 0, 1e-3), ((6, 7, 9, 70), np.float64, INF, 1, 1e-3),
            ((5, 6, 7, 8, 40), np.float32, 0.5, 1, 1e-2), ((1025,), np.float64, 0.0, 1, 1e-5),
            # no level to decompose (L = 0) and the smallest grids with one
            ((3, 3), np.float64, 0.0, 1, 1e-3), ((3, 9), np.float32, 0.0, 0, 1e-3),
            ((4, 5, 3), np.float64, 1.0, 1, 1e-2), ((5, 5), np.float32, 0.0, 1, 1e-7),
            ((3, 3, 3, 3), np.float32, INF, 1, 1e-3)]


@pytest.mark.parametrize("shape,dtype,s,rel,tol", ORTH_E2E)
def test_fused_level_quantize_containers(shape, dtype, s, rel, tol, monkeypatch):
    """mgx_compress on the orthogonal path quantises inside the level kernel
    (no coefficient array exists): containers must still be byte-identical to
    the oracle's, and decode to the oracle's reconstruction."""
    u = smooth_field(shape, dtype)
    o = Oracle("port")
    want, norm = o.compress(u, rel, tol, s)
    comp, tdt = make(shape, dtype)
    eb = mg.REL if rel else mg.ABS
    blob, _ = comp.Compress(torch.from_numpy(u).cuda(), eb, tol, s,
                            norm=norm if rel else 0.0)
    got = blob.cpu().numpy()
    assert got.size == want.size
    assert np.array_equal(np.delete(got, 7), np.delete(want, 7))
    dec = comp.Decompress(blob, eb, tol, s, norm).cpu().numpy()
    assert np.array_equal(dec, o.decompress(want, shape, dtype, rel, tol, s, norm))
    # the decompress side reads the symbols directly when there are no outliers;
    # the dequantise-to-floats route must give the same bits
    monkeypatch.setenv("MGX_NO_SYM_RECOMPOSE", "1")
    comp2, _ = make(shape, dtype)
    monkeypatch.delenv("MGX_NO_SYM_RECOMPOSE")
    dec2 = comp2.Decompress(blob, eb, tol, s, norm).cpu().numpy()
    assert np.array_equal(dec2, dec)


@pytest.mark.parametrize("shape,dtype", [((300, 77), np.float64), ((66, 33, 130), np.float32),
                                         ((5, 6, 7, 8, 40), np.float64)])
def test_single_warp_thomas_kernels(shape, dtype, monkeypatch):
    """Levels with many lines use the single-warp ring kernels instead of the
    producer/consumer pairs; MGX_NO_WS=1 forces them at test sizes."""
    u = smooth_field(shape, dtype)
    monkeypatch.setenv("MGX_NO_WS", "1")
    comp, tdt = make(shape, dtype)
    monkeypatch.delenv("MGX_NO_WS")
    d = torch.from_numpy(u).cuda()
    comp.Decompose(d, 1)
    want = Oracle("port").decompose(u, 1)
    assert np.array_equal(d.cpu().numpy(), want)
    comp.Recompose(d, 1)
    assert np.array_equal(d.cpu().numpy(), Oracle("port").recompose(want, 1))
