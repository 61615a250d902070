"""The drop-in boundary without a GPU: libmgard_b200.so loads, exports every
function include/mgard_b200.h declares, refuses to run without a device (there is
no CPU fallback), and its host-only helpers agree with the oracle."""
import ctypes
import math
import os
import re

import numpy as np
import pytest

import oracle_lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADERS = [os.path.join(ROOT, "include", "mgard_b200.h"),
           os.path.join(ROOT, "include", "mgard_b200_x.h")]


def declared_functions():
    names = []
    for h in HEADERS:
        src = open(h).read()
        src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
        names += re.findall(r"\b(mgx_[a-z0-9_]+)\s*\(", src)
    return sorted(set(names))


def test_header_declares_the_boundary():
    names = declared_functions()
    for must in ("mgx_create", "mgx_destroy", "mgx_plan", "mgx_compress",
                 "mgx_decompress", "mgx_decompose", "mgx_recompose", "mgx_quantize",
                 "mgx_dequantize", "mgx_huff_encode", "mgx_huff_decode",
                 "mgx_compress_host", "mgx_decompress_host", "mgx_x_compress",
                 "mgx_x_decompress", "mgx_metadata_serialize",
                 "mgx_metadata_parse"):
        assert must in names


def test_library_exports_every_declared_symbol():
    from mgard_b200 import _lib
    lib = ctypes.CDLL(_lib.LIB_PATH)
    for name in declared_functions():
        assert hasattr(lib, name), "libmgard_b200.so does not export " + name
    # and the python binding lists exactly the header's functions
    assert sorted(_lib.SIGNATURES) == declared_functions()


def test_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("this checks the GPU-less behaviour")
    import mgard_b200 as mg
    lib = mg.load()
    ctx = ctypes.c_void_p()
    assert lib.mgx_create(ctypes.byref(ctx), 0, None) == 5  # BackendNotAvailable
    with pytest.raises(mg.MgardError) as e:
        mg.Compressor(mg.Hierarchy((9, 9, 9), torch.float32))
    assert e.value.status == 5


def test_product_does_not_import_the_oracle():
    pkg = os.path.join(ROOT, "mgard_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h", ".sh")):
                text = open(os.path.join(dirpath, f)).read()
                assert "mgx_oracle" not in text and "oracle_lib" not in text and \
                    "libmgardx_ref" not in text, f


def test_error_statuses_host_side():
    import torch
    import mgard_b200 as mg
    with pytest.raises(mg.MgardError) as e:
        mg.Hierarchy((2, 2, 2, 2, 2, 2), torch.float32)
    assert e.value.status == 3  # NotSupportHigherNumberOfDimensionsFailure
    with pytest.raises(mg.MgardError) as e:
        mg.Hierarchy((9, 9), torch.float16)
    assert e.value.status == 4  # NotSupportDataTypeFailure
    assert mg.infer_orthogonal_projection(3, float("inf")) is False
    assert mg.infer_orthogonal_projection(4, float("inf")) is True
    assert mg.infer_orthogonal_projection(2, 0.0) is True


@pytest.mark.parametrize("shape", [(5,), (129,), (1000,), (17, 33), (20, 11, 6),
                                   (513, 513, 513), (8, 16, 39, 39, 39969),
                                   (2049, 2049, 2049), (3, 3), (3, 9)])
def test_hierarchy_levels_match_oracle(shape):
    import mgard_b200 as mg
    lib = mg.load()
    D = len(shape)
    shp = (ctypes.c_uint64 * D)(*shape)
    L = ctypes.c_int()
    buf = (ctypes.c_uint64 * (40 * D))()
    assert lib.mgx_hierarchy_levels(D, shp, 2 ** 64 - 1, ctypes.byref(L), buf) == 0
    o = oracle_lib.port()
    buf2 = (ctypes.c_uint64 * (41 * D))()
    L2 = o.mgxo_levels(D, shp, ctypes.c_uint64(2 ** 64 - 1), buf2)
    assert L.value == L2
    n = (L.value + 1) * D
    assert list(buf[:n]) == list(buf2[:n])
    assert list(buf[L.value * D:n]) == list(shape)


def test_finish_norm_and_local_tolerance():
    import mgard_b200 as mg
    lib = mg.load()
    # L-inf: the partial is the norm; L2 with normalised coordinates: sqrt(sum/N)
    assert lib.mgx_finish_norm(0, 1, 3.5, 1000, 1) == 3.5
    assert lib.mgx_finish_norm(1, 0, 250.0, 1000, 1) == math.sqrt(250.0 / 1000)
    inf = float("inf")
    f32 = np.float32
    assert lib.mgx_local_abs_tol(0, 1, 1.5, 1e-3, inf, 8) == float(f32(1e-3) * f32(1.5))
    assert lib.mgx_local_abs_tol(1, 0, 1.5, 1e-3, inf, 8) == 1e-3
    assert lib.mgx_local_abs_tol(1, 0, 1.5, 1e-3, 0.0, 4) == math.sqrt(1e-3 * 1e-3 / 4)
    assert lib.mgx_local_abs_tol(1, 1, 2.0, 1e-3, 0.0, 4) == \
        math.sqrt((1e-3 * 2.0) * (1e-3 * 2.0) / 4)


@pytest.mark.parametrize("shape,devices,want", [
    ((2049, 2049, 2049), 8, (0, 257, 8)),   # BASELINE cfg 5
    ((2049, 2049, 2049), 1, (0, 2049, 1)),
    ((8, 16, 39, 39, 39969), 2, (4, 19985, 2)),
    ((100, 300, 200), 4, (1, 75, 4)),
    ((5, 5), 8, (0, 1, 5)),
])
def test_max_dim_decomposition(shape, devices, want):
    from mgard_b200.sharding import DomainDecomposer
    dd = DomainDecomposer(shape, devices)
    assert (dd.dim, dd.size, dd.num_subdomains) == want
    total = 0
    for i in range(dd.num_subdomains):
        shp, off = dd.subdomain_shape(i)
        assert off == i * dd.size
        assert all(shp[d] == shape[d] for d in range(len(shape)) if d != dd.dim)
        total += shp[dd.dim]
    assert total == shape[dd.dim]
    with pytest.raises(IndexError):
        dd.subdomain_shape(dd.num_subdomains)


def test_libzstd_is_found_at_run_time_and_the_config_carries_its_level():
    """lossless_type::Huffman_Zstd binds libzstd.so.1 with dlopen (this image ships the
    library without headers); Config::zstd_compress_level defaults to 3 (Config.cpp:44)"""
    from mgard_b200 import _lib, highlevel as hl
    lib = _lib.load()
    assert lib.mgx_x_have_zstd() in (0, 1)
    if os.path.exists("/usr/lib/x86_64-linux-gnu/libzstd.so.1"):
        assert lib.mgx_x_have_zstd() == 1
    cfg = hl.Config()
    assert cfg.zstd_compress_level == 3 and cfg.lossless == 0


@pytest.mark.parametrize("n,chunk", [(1, 32768), (32768, 32768), (32769, 32768),
                                     (200001, 4096), (70000, 65536)])
def test_lz4_container_round_trip_and_checks(n, chunk):
    """mgx_x_lz4_wrap / mgx_x_lz4_unwrap on host buffers (the reference's portable LZ4
    container, Lossless/LZ4/LZ4.hpp:28-36): layout, round trip, corrupt input refused"""
    from mgard_b200 import _lib
    lib = _lib.load()
    if not lib.mgx_x_have_lz4():
        pytest.skip("no liblz4.so.1 on this machine")
    rng = np.random.default_rng(n)
    # half compressible, half noise
    data = np.where(rng.random(n) < 0.5, 7, rng.integers(0, 256, n)).astype(np.uint8)
    nb = ctypes.c_uint64()
    assert lib.mgx_x_lz4_wrap(data.ctypes.data, n, chunk, None, 0, ctypes.byref(nb)) == 2
    box = np.zeros(nb.value, np.uint8)
    assert lib.mgx_x_lz4_wrap(data.ctypes.data, n, chunk, box.ctypes.data, box.size,
                              ctypes.byref(nb)) == 0 and nb.value == box.size
    nchunks = (n - 1) // chunk + 1
    assert box[:8].tobytes() == b"MGXLZ4P\x00"
    assert [int(v) for v in box[8:32].view(np.uint64)] == [n, chunk, nchunks]
    sizes = box[32:32 + 8 * nchunks].view(np.uint64)
    packed = int(box[32 + 8 * nchunks:40 + 8 * nchunks].view(np.uint64)[0])
    assert int(sizes.sum()) == packed == box.size - (40 + 8 * nchunks)
    out = np.zeros(n, np.uint8)
    assert lib.mgx_x_lz4_unwrap(box.ctypes.data, box.size, out.ctypes.data, n,
                                ctypes.byref(nb)) == 0 and nb.value == n
    assert np.array_equal(out, data)
    # corrupt: signature, a chunk size, truncation, announced size
    for mutate in ("sig", "chunk", "trunc", "n"):
        bad = box.copy()
        if mutate == "sig":
            bad[3] ^= 1
        elif mutate == "chunk":
            bad[32:40] = np.array([int(sizes[0]) + 3], np.uint64).view(np.uint8)
        elif mutate == "trunc":
            bad = bad[:-2].copy()
        else:
            bad[8:16] = np.array([n + chunk + 1], np.uint64).view(np.uint8)
        assert lib.mgx_x_lz4_unwrap(bad.ctypes.data, bad.size, out.ctypes.data, n,
                                    ctypes.byref(nb)) != 0, mutate
