"""The C restatement against the reference itself (oracle/_ref, compiled from
/root/reference by oracle/Makefile) on seeded inputs.  Skipped where the
reference library has not been built (it cannot be built on the GPU box)."""
import numpy as np
import pytest

import oracle_lib
from oracle_lib import Oracle, smooth_field

pytestmark = pytest.mark.skipif(not oracle_lib.have_ref(),
                                reason="oracle/_ref not built (needs /root/reference)")
INF = float("inf")

SHAPES = [(5,), (64,), (129,), (9, 9), (17, 33), (40, 31), (9, 9, 9), (20, 11, 6),
          (33, 17, 65), (5, 4, 9, 7), (3, 5, 5, 9, 9)]


@pytest.mark.parametrize("shape", SHAPES)
@pytest.mark.parametrize("dtype", [np.float32, np.float64])
@pytest.mark.parametrize("orth", [0, 1])
def test_decompose_recompose_bit_exact(shape, dtype, orth):
    if len(shape) > 3 and not orth:
        pytest.skip("D > 3 always runs with the orthogonal correction "
                    "(infer_orthogonal_projection, Compressor.h:38-40)")
    u = smooth_field(shape, dtype, seed=3)
    r, p = Oracle("ref"), Oracle("port")
    cr, cp = r.decompose(u, orth), p.decompose(u, orth)
    assert np.array_equal(cr, cp)
    assert np.array_equal(r.recompose(cr, orth), p.recompose(cp, orth))


@pytest.mark.parametrize("shape", [(129,), (17, 33), (20, 11, 6), (5, 4, 9, 7)])
@pytest.mark.parametrize("dtype", [np.float32, np.float64])
@pytest.mark.parametrize("s,rel,tol", [(INF, 1, 1e-3), (0.0, 1, 1e-4), (1.0, 0, 1e-3),
                                       (INF, 0, 1e-6)])
def test_quantize_dequantize_bit_exact(shape, dtype, s, rel, tol):
    u = smooth_field(shape, dtype, seed=5)
    r, p = Oracle("ref"), Oracle("port")
    orth = not (np.isinf(s) and len(shape) <= 3)
    c = p.decompose(u, orth)
    norm = 1.2345
    qr, qp = r.quantize(c, rel, tol, s, norm), p.quantize(c, rel, tol, s, norm)
    assert np.array_equal(qr, qp)
    assert np.array_equal(r.dequantize(qr, dtype, rel, tol, s, norm),
                          p.dequantize(qp, dtype, rel, tol, s, norm))


@pytest.mark.parametrize("seed", range(12))
def test_huffman_container(seed):
    rng = np.random.default_rng(seed)
    n = int(rng.integers(1, 60000))
    scale = rng.uniform(0.3, 300)
    codes = (rng.laplace(0, scale, n).astype(np.int64) + 4096)
    if seed % 3 == 0:  # force outliers on both sides
        codes[rng.integers(0, n, 5)] = rng.integers(-10 ** 9, 10 ** 9, 5)
    got = oracle_lib.fresh("ref", "huffman_compress", codes)
    allowed = oracle_lib.ref_allowed_outputs("huffman_compress", codes)
    assert any(oracle_lib.same_container(got, a) for a in allowed)
    r, p = Oracle("ref"), Oracle("port")
    # each decodes the other's stream
    assert np.array_equal(p.huffman_decompress(got, n), codes)
    assert np.array_equal(r.huffman_decompress(allowed[0], n), codes)


@pytest.mark.parametrize("shape,dtype,s,rel,tol", [
    ((33, 33, 33), np.float32, INF, 1, 1e-3), ((65, 65), np.float64, 0.0, 1, 1e-4),
    ((257,), np.float32, INF, 0, 1e-4), ((4, 6, 9, 9, 17), np.float64, 1.0, 1, 1e-3),
    ((40, 30, 21), np.float32, INF, 1, 1e-2)])
def test_compress_decompress(shape, dtype, s, rel, tol):
    u = smooth_field(shape, dtype, seed=9)
    got, norm = oracle_lib.fresh("ref", "compress", u, rel, tol, s)
    allowed = oracle_lib.ref_allowed_outputs("compress", u, rel, tol, s)
    assert norm == allowed[0][1]
    assert any(oracle_lib.same_container(got, a[0]) for a in allowed)
    r, p = Oracle("ref"), Oracle("port")
    dr = r.decompress(allowed[0][0], shape, dtype, rel, tol, s, norm)
    dp = p.decompress(got, shape, dtype, rel, tol, s, norm)
    assert np.array_equal(dr, dp)
    if np.isinf(s):
        bound = tol * (np.abs(u).max() if rel else 1.0)
        assert np.abs(dp.astype(np.float64) - u).max() <= bound


def test_openmp_backend_matches_serial_size():
    """reference OPENMP path = the CPU baseline bench.py times"""
    u = smooth_field((33, 33, 33), np.float32)
    r = Oracle("ref")
    a, _ = r.compress(u, 1, 1e-3, INF, backend=0)
    b, _ = r.compress(u, 1, 1e-3, INF, backend=1)
    assert a.size == b.size


@pytest.mark.parametrize("shape,dtype,s,rel,tol", [
    ((33, 40, 65), np.float32, float("inf"), 1, 1e-3),
    ((129, 70), np.float64, 0.0, 0, 1e-4),
])
def test_reference_huffman_zstd_record_is_a_size_and_a_zstd_frame(shape, dtype, s, rel, tol):
    """lossless_type::Huffman_Zstd run by the REFERENCE's own writer and reader
    (Lossless.hpp:188-192, Zstd.hpp:72-141; its ZSTD_* calls forwarded to the machine's
    libzstd.so.1 by oracle/ref_wrap_huff.cpp): the record is [u64 n][ZSTD frame], the
    frame -- read with another zstd implementation -- is the n bytes of the plain
    Huffman container, and the reference decodes it to the same field.  This is the
    layout mgx_highlevel.cpp writes and reads (tests/test_highlevel_gpu.py)."""
    pa = pytest.importorskip("pyarrow")
    o = Oracle("ref")
    if not o.lib.ref_have_zstd():
        pytest.skip("no libzstd.so.1 on this machine")
    u = smooth_field(shape, dtype)
    plain, norm = o.compress(u, rel, tol, s)
    back_plain = o.decompress(plain, shape, dtype, rel, tol, s, norm)
    o.lib.ref_set_lossless(2)
    try:
        rec, norm_z = o.compress(u, rel, tol, s)
        back = o.decompress(rec, shape, dtype, rel, tol, s, norm_z)
    finally:
        o.lib.ref_set_lossless(0)
    assert norm_z == norm
    n = int(rec[:8].view(np.uint64)[0])
    assert n == plain.size
    assert rec[8:12].tobytes() == bytes.fromhex("28b52ffd")   # zstd frame magic
    body = np.frombuffer(pa.Codec("zstd").decompress(rec[8:].tobytes(), decompressed_size=n,
                                                     asbytes=True), np.uint8)
    # (byte 7 of the Huffman container is a pad byte the reference leaves uninitialised)
    assert np.array_equal(np.delete(body, 7), np.delete(plain, 7))
    assert np.array_equal(back, back_plain)


@pytest.mark.parametrize("shape,dtype,s,rel,tol", [
    ((33, 40, 65), np.float32, float("inf"), 1, 1e-3),
    ((129, 70), np.float64, 0.0, 0, 1e-4),
])
def test_reference_huffman_lz4_container_both_ways(shape, dtype, s, rel, tol):
    """lossless_type::Huffman_LZ4 = the reference's portable LZ4 container
    (Lossless/LZ4/LZ4.hpp:28-36, LZ4Kernels.hpp:202-262) around the serialised Huffman
    container.  The library's host reader (liblz4 blocks, mgx_x_lz4_unwrap) decodes the
    REFERENCE's container to its plain Huffman container, and the REFERENCE decodes the
    container the library's writer (mgx_x_lz4_wrap) makes of that Huffman container to
    the same field; the block format is shared, the match finder is not (sizes differ)."""
    import ctypes
    from mgard_b200 import _lib
    my = _lib.load()
    if not my.mgx_x_have_lz4():
        pytest.skip("no liblz4.so.1 on this machine")
    o = Oracle("ref")
    u = smooth_field(shape, dtype)
    plain, norm = o.compress(u, rel, tol, s)
    want = o.decompress(plain, shape, dtype, rel, tol, s, norm)
    nb = ctypes.c_uint64()
    o.lib.ref_set_lossless(1)
    try:
        rec, _ = o.compress(u, rel, tol, s)
        assert rec[:8].tobytes() == b"MGXLZ4P\x00"
        n, chunk, nchunks = (int(v) for v in rec[8:32].view(np.uint64))
        assert n == plain.size and chunk == 1 << 15 and nchunks == (n - 1) // chunk + 1
        out = np.zeros(n, np.uint8)
        assert my.mgx_x_lz4_unwrap(rec.ctypes.data, rec.size, out.ctypes.data, out.size,
                                   ctypes.byref(nb)) == 0 and nb.value == n
        # (byte 7 of the Huffman container is a pad byte the reference leaves uninitialised)
        assert np.array_equal(np.delete(out, 7), np.delete(plain, 7))
        mine = np.zeros(2 * n + 4096, np.uint8)
        assert my.mgx_x_lz4_wrap(plain.ctypes.data, n, chunk, mine.ctypes.data, mine.size,
                                 ctypes.byref(nb)) == 0
        mine = mine[:nb.value].copy()
        assert np.array_equal(o.decompress(mine, shape, dtype, rel, tol, s, norm), want)
        assert np.array_equal(o.decompress(rec, shape, dtype, rel, tol, s, norm), want)
    finally:
        o.lib.ref_set_lossless(0)
