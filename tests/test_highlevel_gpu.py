"""GPU tests of the high-level path (mgard_x::compress / decompress, reference
include/compress_x.hpp:31-178) through the C ABI of include/mgard_b200_x.h.

The stream is header + [u64 size][payload] per subdomain.  Every payload must be
the byte-exact container the CPU oracle writes for that subdomain with the
tolerance the reference derives for it (calc_local_abs_tol, global norm), the
header must parse, and decompress must honour the bound on every element."""
import ctypes
import math
import os
import struct
import subprocess

import numpy as np
import pytest
import torch

import oracle_lib
from oracle_lib import Oracle, smooth_field, same_container

import mgard_b200 as mg
from mgard_b200 import _lib, highlevel as hl

pytestmark = pytest.mark.gpu
INF = float("inf")
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def records(stream):
    """-> (metadata, [payload bytes per subdomain])"""
    stream = np.asarray(stream.cpu() if hasattr(stream, "cpu") else stream)
    md, pos = hl.read_metadata(bytes(stream[:4096]))
    out = []
    while pos < stream.size:
        (n,) = struct.unpack_from("<Q", stream, pos)
        pos += 8
        out.append(stream[pos:pos + n].copy())
        pos += n
    assert pos == stream.size
    return md, out


def footprint(shape, dtype, s, prefetch=0):
    shp = (ctypes.c_uint64 * len(shape))(*shape)
    return _lib.load().mgx_estimate_footprint(len(shape), shp,
                                              1 if dtype == np.float64 else 0,
                                              s, prefetch)


def budget_for_halvings(shape, dtype, s, halvings):
    """a max_memory_footprint that makes the library halve the largest dimension
    at least once and at most `halvings` times"""
    big = max(shape)
    size = big
    for _ in range(halvings):
        size = (size - 1) // 2 + 1
    sub = list(shape)
    sub[shape.index(big)] = size
    avail = min(footprint(sub, dtype, s, 1) + 1, footprint(shape, dtype, s))
    return hl.Config(max_memory_footprint=avail), size


def oracle_container(u, rel, tol, s, norm):
    """the oracle's container for u with a GIVEN norm (its whole-path entry
    recomputes the norm; a parallel fp64 sum differs from its serial one in the
    last bits, exactly as the reference's CUDA and SERIAL back ends differ)"""
    o = Oracle("port")
    orth = mg.infer_orthogonal_projection(u.ndim, s)
    codes = o.quantize(o.decompose(u, orth), rel, tol, s, norm)
    return o.huffman_compress(codes)


def slabs_of(md):
    shape = [md.shape[d] for d in range(md.D)]
    if not md.decomposed:
        return [tuple(slice(None) for _ in shape)]
    out = []
    n, size, dim = shape[md.dd_dim], md.dd_size, md.dd_dim
    for i in range((n - 1) // size + 1):
        sl = [slice(None)] * md.D
        sl[dim] = slice(i * size, min((i + 1) * size, n))
        out.append(tuple(sl))
    return out


def check_bound(u, back, rel, tol, s):
    u64, b64 = u.astype(np.float64), np.asarray(back).astype(np.float64)
    if math.isinf(s):
        scale = np.abs(u64).max() if rel else 1.0
        assert np.abs(u64 - b64).max() <= tol * scale
    else:
        # Utilities/ErrorCalculator.h:56-97 with normalised coordinates
        err = math.sqrt(((u64 - b64) ** 2).sum() / u64.size)
        scale = math.sqrt((u64 ** 2).sum() / u64.size) if rel else 1.0
        assert err <= tol * scale


@pytest.mark.parametrize("shape,dtype,s,mode", [
    ((33, 40, 65), np.float32, INF, hl.REL),
    ((129, 70), np.float64, 0.0, hl.ABS),
    ((5000,), np.float32, INF, hl.ABS),          # smaller than its decodebook
    ((100001,), np.float32, INF, hl.ABS),
    ((9, 10, 12, 60), np.float64, 1.0, hl.REL),
])
@pytest.mark.parametrize("where", ["device", "host"])
def test_single_domain_stream(shape, dtype, s, mode, where):
    u = smooth_field(shape, dtype)
    tol = 1e-3
    x = torch.from_numpy(u).cuda() if where == "device" else u
    stream = hl.compress(x, tol, s, mode)
    assert (where == "device") == hasattr(stream, "is_cuda")
    md, recs = records(stream)
    assert not md.decomposed and md.dd_size == shape[0] and len(recs) == 1
    assert md.backend == 3 and md.dtype == (1 if dtype == np.float64 else 0)
    tolT = float(np.float32(tol)) if dtype == np.float32 else tol
    assert md.tol == tolT and md.mode == mode
    want, norm = Oracle("port").compress(u, int(mode == hl.REL), tolT, s)
    if mode == hl.REL and not math.isinf(s):
        assert md.norm == pytest.approx(norm, rel=1e-12)
        norm = md.norm
        want = oracle_container(u, 1, tolT, s, norm)
    if want.size > u.nbytes:
        # CR < 1 (the 64 KiB decodebook outweighs a tiny input): the reference
        # stores the subdomain itself (GPUPipelines.hpp:139-156)
        assert np.array_equal(recs[0].view(dtype).reshape(shape), u)
    else:
        assert same_container(recs[0], want)
    if mode == hl.REL:
        assert md.norm == norm
    back = hl.decompress(stream)
    assert tuple(back.shape) == shape
    check_bound(u, back.cpu().numpy() if where == "device" else back,
                mode == hl.REL, tol, s)


def test_host_and_device_streams_are_identical():
    u = smooth_field((40, 33, 50), np.float32)
    a = hl.compress(u, 1e-3, INF, hl.REL)
    b = hl.compress(torch.from_numpy(u).cuda(), 1e-3, INF, hl.REL).cpu().numpy()
    assert np.array_equal(a, b)
    # a host stream decodes on the device path and vice versa
    x = hl.decompress(torch.from_numpy(a).cuda()).cpu().numpy()
    y = hl.decompress(b)
    assert np.array_equal(x, y)


@pytest.mark.parametrize("shape,dtype,s,mode", [
    ((140, 80, 66), np.float32, INF, hl.REL),    # dim 0: contiguous slabs
    ((40, 180, 66), np.float32, INF, hl.ABS),    # dim 1: pitched copies
    ((24, 60, 190), np.float64, INF, hl.REL),    # dim 2
    ((600, 257), np.float64, 0.0, hl.REL),       # L2: tolerance split over slabs
    ((12, 10, 14, 120), np.float64, 1.0, hl.ABS),  # 4-D, s-norm
    ((64, 40, 33), np.float32, INF, hl.REL),     # slabs too small to compress
])
@pytest.mark.parametrize("where", ["device", "host"])
def test_max_dim_decomposition(shape, dtype, s, mode, where):
    """max_memory_footprint forces MaxDim slabs (DomainDecomposer.hpp:192-223);
    every record is the oracle's container for that slab compressed with an ABS
    bound of calc_local_abs_tol (ErrorToleranceCalculator.hpp:126-147)."""
    u = smooth_field(shape, dtype)
    tol = 1e-3
    cfg, size = budget_for_halvings(shape, dtype, s, 3)
    big = max(shape)
    x = torch.from_numpy(u).cuda() if where == "device" else u
    stream = hl.compress(x, tol, s, mode, config=cfg)
    md, recs = records(stream)
    assert md.decomposed and md.dd_type == 0
    assert md.dd_dim == shape.index(big) and size <= md.dd_size < big
    sls = slabs_of(md)
    assert len(recs) == len(sls) and len(recs) >= 2
    # global norm, local tolerance
    lib = _lib.load()
    code = 1 if dtype == np.float64 else 0
    tolT = float(np.float32(tol)) if dtype == np.float32 else tol
    rel = int(mode == hl.REL)
    norm = 1.0
    if rel:
        u64 = u.astype(np.float64)
        part = np.abs(u64).max() if math.isinf(s) else (u64 ** 2).sum()
        norm = lib.mgx_finish_norm(code, int(math.isinf(s)), part, u.size, 1)
        assert md.norm == pytest.approx(norm, rel=1e-6)
        norm = md.norm
    ltol = lib.mgx_local_abs_tol(code, rel, norm, tolT, s, len(sls))
    port = Oracle("port")
    for rec, sl in zip(recs, sls):
        piece = np.ascontiguousarray(u[sl])
        want, _ = port.compress(piece, 0, ltol, s)
        if want.size > piece.nbytes:
            assert np.array_equal(rec.view(dtype).reshape(piece.shape), piece)
        else:
            assert same_container(rec, want)
    back = hl.decompress(stream)
    check_bound(u, back.cpu().numpy() if where == "device" else back, rel, tol, s)


@pytest.mark.parametrize("shape,dtype,s,mode,decomposed", [
    ((33, 40, 65), np.float32, INF, hl.REL, False),
    ((129, 70), np.float64, 0.0, hl.ABS, False),
    ((24, 60, 190), np.float64, INF, hl.REL, True),     # slabs along dim 2: strided copies back
    ((600, 257), np.float32, 0.0, hl.REL, True),        # slabs along dim 0
])
@pytest.mark.parametrize("where", ["device", "host"])
def test_compress_with_dryrun_leaves_the_reconstruction_in_the_input(shape, dtype, s, mode,
                                                                     decomposed, where):
    """Config::compress_with_dryrun (Compressor.hpp:239-242: Dequantize + Recompose after
    Serialize; GPUPipelines.hpp:198-202: SubdomainToOriginal): after the call the INPUT
    array holds what decompress returns for the stream, bit for bit; the stream itself is
    the one a plain compress writes."""
    u = smooth_field(shape, dtype)
    tol = 1e-3
    cfg = budget_for_halvings(shape, dtype, s, 3)[0] if decomposed else hl.Config()
    plain = hl.compress(torch.from_numpy(u).cuda() if where == "device" else u.copy(),
                        tol, s, mode, config=cfg)
    cfg.compress_with_dryrun = 1
    x = torch.from_numpy(u).cuda() if where == "device" else u.copy()
    stream = hl.compress(x, tol, s, mode, config=cfg)
    as_np = lambda t: t.cpu().numpy() if hasattr(t, "cpu") else np.asarray(t)
    assert np.array_equal(as_np(stream), as_np(plain))
    md, recs = records(stream)
    assert bool(md.decomposed) == decomposed
    back = as_np(hl.decompress(stream))
    got = as_np(x)
    assert not np.array_equal(got, u), "the input was not touched"
    assert np.array_equal(got.view(np.uint8), back.view(np.uint8))
    check_bound(u, got, int(mode == hl.REL), tol, s)


def test_dryrun_is_refused_by_the_multi_gpu_entry():
    u = smooth_field((40, 30, 50), np.float32)
    with pytest.raises(hl.MgardError):
        hl.compress_multi(u, 1e-3, INF, hl.REL, [0],
                          config=hl.Config(compress_with_dryrun=1))


def _restream(stream, payloads):
    """the stream's header followed by the given record payloads"""
    stream = np.asarray(stream)
    _, pos = hl.read_metadata(bytes(stream[:4096]))
    parts = [stream[:pos].tobytes()]
    for p in payloads:
        parts.append(struct.pack("<Q", len(p)))
        parts.append(bytes(p))
    return np.frombuffer(b"".join(parts), np.uint8).copy()


@pytest.mark.parametrize("shape,dtype,s,mode,decomposed", [
    ((33, 40, 65), np.float32, INF, hl.REL, False),
    ((257, 130), np.float64, 0.0, hl.ABS, False),
    ((24, 60, 190), np.float64, INF, hl.REL, True),
])
@pytest.mark.parametrize("where", ["device", "host"])
def test_huffman_zstd_second_stage(shape, dtype, s, mode, decomposed, where):
    """lossless_type::Huffman_Zstd (Lossless.hpp:188-192, 250-253; Zstd.hpp:72-141): every
    record is [u64 n][ZSTD frame of the n bytes of the Huffman container].  The frames
    are read back with ANOTHER zstd implementation (pyarrow's) and must hold the records
    of the plain Huffman stream; frames written by that other implementation at level 19
    and spliced into the stream decode to the same field."""
    pa = pytest.importorskip("pyarrow")
    if not _lib.load().mgx_x_have_zstd():
        pytest.skip("no libzstd.so.1 on this machine")
    u = smooth_field(shape, dtype)
    tol = 1e-3
    cfg = budget_for_halvings(shape, dtype, s, 3)[0] if decomposed else hl.Config()
    as_np = lambda t: t.cpu().numpy() if hasattr(t, "cpu") else np.asarray(t)
    x = torch.from_numpy(u).cuda() if where == "device" else u
    plain = as_np(hl.compress(x, tol, s, mode, config=cfg))
    cfg.lossless = hl.HUFFMAN_ZSTD
    stream = as_np(hl.compress(x, tol, s, mode, config=cfg))
    md0, recs0 = records(plain)
    md, recs = records(stream)
    assert md0.lossless == 0 and md.lossless == 2 and bool(md.decomposed) == decomposed
    assert len(recs) == len(recs0)
    codec = pa.Codec("zstd", compression_level=19)
    other = []
    for r, r0 in zip(recs, recs0):
        n = int(r[:8].view(np.uint64)[0])
        assert n == r0.size and r.size < r0.size + 64
        body = codec.decompress(r[8:].tobytes(), decompressed_size=n, asbytes=True)
        assert body == r0.tobytes(), "the frame does not hold the Huffman container"
        other.append(struct.pack("<Q", n) + codec.compress(r0.tobytes(), asbytes=True))
    want = as_np(hl.decompress(plain))
    got = as_np(hl.decompress(torch.from_numpy(stream).cuda() if where == "device" else stream))
    assert np.array_equal(got.view(np.uint8), want.view(np.uint8))
    foreign = _restream(stream, other)
    got2 = as_np(hl.decompress(foreign))
    assert np.array_equal(got2.view(np.uint8), want.view(np.uint8))
    check_bound(u, got, int(mode == hl.REL), tol, s)
    # a truncated frame is refused
    bad = _restream(stream, [bytes(r[:-5]) for r in recs])
    with pytest.raises(hl.MgardError):
        hl.decompress(bad)


@pytest.mark.xfail(strict=False, reason="written after the round's GPU minutes were spent: "
                   "the host stage is pinned against the reference on the CPU "
                   "(tests/test_oracle_vs_ref.py, tests/test_cabi.py) and the pipeline code "
                   "is the Zstd one's, but this test has not run on a GPU yet")
@pytest.mark.parametrize("shape,dtype,s,mode,decomposed", [
    ((33, 40, 65), np.float32, INF, hl.REL, False),
    ((24, 60, 190), np.float64, INF, hl.REL, True),
])
@pytest.mark.parametrize("where", ["device", "host"])
def test_huffman_lz4_second_stage(shape, dtype, s, mode, decomposed, where):
    """lossless_type::Huffman_LZ4 (Lossless.hpp:182-186, LZ4/LZ4.hpp:28-36): every record is
    the reference's portable LZ4 container of the Huffman container; the records unwrap
    to those of the plain Huffman stream and the stream decodes to the same field."""
    lib = _lib.load()
    if not lib.mgx_x_have_lz4():
        pytest.skip("no liblz4.so.1 on this machine")
    u = smooth_field(shape, dtype)
    tol = 1e-3
    cfg = budget_for_halvings(shape, dtype, s, 3)[0] if decomposed else hl.Config()
    as_np = lambda t: t.cpu().numpy() if hasattr(t, "cpu") else np.asarray(t)
    x = torch.from_numpy(u).cuda() if where == "device" else u
    plain = as_np(hl.compress(x, tol, s, mode, config=cfg))
    cfg.lossless = hl.HUFFMAN_LZ4
    stream = as_np(hl.compress(x, tol, s, mode, config=cfg))
    md0, recs0 = records(plain)
    md, recs = records(stream)
    assert md.lossless == hl.HUFFMAN_LZ4 and len(recs) == len(recs0)
    nb = ctypes.c_uint64()
    for r, r0 in zip(recs, recs0):
        r = np.ascontiguousarray(r)
        assert r[:8].tobytes() == b"MGXLZ4P\x00"
        out = np.zeros(r0.size, np.uint8)
        assert lib.mgx_x_lz4_unwrap(r.ctypes.data, r.size, out.ctypes.data, out.size,
                                    ctypes.byref(nb)) == 0 and nb.value == r0.size
        assert np.array_equal(out, r0)
    want = as_np(hl.decompress(plain))
    got = as_np(hl.decompress(torch.from_numpy(stream).cuda() if where == "device" else stream))
    assert np.array_equal(got.view(np.uint8), want.view(np.uint8))
    check_bound(u, got, int(mode == hl.REL), tol, s)
    bad = _restream(stream, [bytes(r[:-5]) for r in recs])
    with pytest.raises(hl.MgardError):
        hl.decompress(bad)


def test_raw_fallback_when_a_subdomain_does_not_compress():
    """CR < 1 -> the record is the raw subdomain (GPUPipelines.hpp:139-156) and
    decompress copies it back (:407-421)."""
    rng = np.random.default_rng(3)
    u = rng.standard_normal((40, 30, 20)).astype(np.float32)
    for x in (u, torch.from_numpy(u).cuda()):
        stream = hl.compress(x, 1e-9, INF, hl.REL)
        md, recs = records(stream)
        assert len(recs) == 1 and recs[0].size == u.nbytes
        assert np.array_equal(recs[0].view(np.float32).reshape(u.shape), u)
        back = hl.decompress(stream)
        back = back.cpu().numpy() if hasattr(back, "cpu") else back
        assert np.array_equal(back, u)


def test_status_codes():
    lib = _lib.load()
    cfg = hl.Config()
    u = smooth_field((17, 17, 17), np.float32)
    shp = (ctypes.c_uint64 * 6)(17, 17, 17, 1, 1, 1)
    out = np.empty(1 << 20, np.uint8)
    p, size = ctypes.c_void_p(out.ctypes.data), ctypes.c_size_t(out.size)

    def call(D=3, dtype=0, cfg=cfg, cap=out.size):
        size.value = cap
        return lib.mgx_x_compress(D, dtype, shp, 1e-3, INF, 0, u.ctypes.data,
                                  ctypes.byref(p), ctypes.byref(size),
                                  ctypes.byref(cfg), 1)
    assert call() == 0
    good = out[:size.value].copy()
    assert call(D=6) == 3           # NotSupportHigherNumberOfDimensionsFailure
    assert call(dtype=7) == 4       # NotSupportDataTypeFailure
    assert call(cfg=hl.Config(dev_type=hl.OPENMP)) == 5   # BackendNotAvailable
    assert call(cfg=hl.Config(lossless=hl.CPU_LOSSLESS)) == 1  # not built: Failure
    assert call(cap=64) == 2        # OutputTooLargeFailure
    assert call(cap=300) == 2
    assert b"too large" in lib.mgx_x_last_error()
    # decompress: truncated / corrupted streams fail, never crash
    for bad in (good[:40], good[:good.size // 2]):
        with pytest.raises(mg.MgardError):
            hl.decompress(bad)
    # library-allocated output (output_pre_allocated = false), host input
    q, n = ctypes.c_void_p(), ctypes.c_size_t()
    assert lib.mgx_x_compress(3, 0, shp, 1e-3, INF, 0, u.ctypes.data,
                              ctypes.byref(q), ctypes.byref(n),
                              ctypes.byref(cfg), 0) == 0
    got = np.ctypeslib.as_array(ctypes.cast(q, ctypes.POINTER(ctypes.c_uint8)),
                                (n.value,)).copy()
    assert np.array_equal(got, good)
    r = ctypes.c_void_p()
    shape_out = (ctypes.c_uint64 * 5)()
    D, dt = ctypes.c_int(), ctypes.c_int()
    assert lib.mgx_x_decompress(q, n.value, ctypes.byref(r), shape_out,
                                ctypes.byref(D), ctypes.byref(dt),
                                ctypes.byref(cfg), 0) == 0
    assert D.value == 3 and dt.value == 0 and list(shape_out[:3]) == [17, 17, 17]
    back = np.ctypeslib.as_array(ctypes.cast(r, ctypes.POINTER(ctypes.c_float)),
                                 (u.size,)).reshape(u.shape)
    assert np.abs(back - u).max() <= 1e-3 * np.abs(u).max()
    libc = ctypes.CDLL(None)
    libc.free.argtypes = [ctypes.c_void_p]
    libc.free(q)
    libc.free(r)
    assert lib.mgx_x_release_cache(ctypes.byref(cfg)) == 0
    # the cache is rebuilt on demand
    assert call() == 0 and np.array_equal(out[:size.value], good)


def test_pin_memory_api():
    lib = _lib.load()
    cfg = hl.Config()
    buf = np.zeros(1 << 20, np.uint8)
    assert lib.mgx_x_check_memory_pinned(buf.ctypes.data, ctypes.byref(cfg)) == 0
    assert lib.mgx_x_pin_memory(buf.ctypes.data, buf.size, ctypes.byref(cfg)) == 0
    assert lib.mgx_x_check_memory_pinned(buf.ctypes.data, ctypes.byref(cfg)) == 1
    assert lib.mgx_x_unpin_memory(buf.ctypes.data, ctypes.byref(cfg)) == 0
    assert lib.mgx_x_check_memory_pinned(buf.ctypes.data, ctypes.byref(cfg)) == 0


def test_reference_decoder_reads_a_decomposed_record():
    """interoperability: the reference's own low-level Decompress (oracle/_ref)
    decodes a record taken out of our decomposed stream."""
    if not oracle_lib.have_ref():
        pytest.skip("oracle/_ref was not built")
    shape = (140, 80, 66)
    u = smooth_field(shape, np.float32)
    cfg, _ = budget_for_halvings(shape, np.float32, INF, 2)
    stream = hl.compress(u, 1e-3, INF, hl.REL, config=cfg)
    md, recs = records(stream)
    sls = slabs_of(md)
    assert md.decomposed and len(recs) == len(sls) >= 2
    ltol = _lib.load().mgx_local_abs_tol(0, 1, md.norm, md.tol, INF, len(sls))
    for rec, sl in zip(recs, sls):
        piece = u[sl]
        back = oracle_lib.fresh("ref", "decompress", rec, piece.shape, np.float32,
                                0, ltol, INF, 1.0)
        assert np.abs(back - piece).max() <= 1e-3 * np.abs(u).max()


def test_cpp_api_example_program():
    """a program written against the reference's C++ API (namespace mgard_x,
    examples/mgard-x/HighLevelAPIs/Example.cpp:18-36) built with
    include/mgard_b200/compress_x.hpp"""
    exe = os.path.join(ROOT, "tests", "cpp", "_build", "highlevel_example")
    if not os.path.exists(exe):
        pytest.fail("tests/cpp/_build/highlevel_example missing: run "
                    "__graft_entry__.build()")
    r = subprocess.run([exe], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "ALL OK" in r.stdout
