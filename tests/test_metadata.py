"""The "MGARD" header (reference src/mgard-x/Metadata/Metadata.cpp:259-762,
src/mgard.proto) without a GPU: mgx_metadata_serialize must reproduce, byte for
byte, the headers tests/golden/make_header_golden.py made with the official
protobuf runtime from the reference's own .proto; mgx_metadata_parse must read
them back and apply the reference's checks."""
import ctypes
import json
import math
import os
import struct
import zlib

import numpy as np
import pytest

from mgard_b200 import _lib

HERE = os.path.dirname(os.path.abspath(__file__))
GOLDEN = json.load(open(os.path.join(HERE, "golden", "header_v1.json")))


def make_md(c):
    md = _lib.MgxMetadata()
    md.version[0], md.version[1], md.version[2] = 1, 0, 0
    md.D = len(c["shape"])
    for d, n in enumerate(c["shape"]):
        md.shape[d] = n
    md.dtype = 1 if c["dtype"] == "f64" else 0
    md.mode = 0 if c["mode"] == "REL" else 1
    s = math.inf if c["s"] == "inf" else float(c["s"])
    md.linf = 1 if math.isinf(s) else 0
    md.s, md.norm, md.tol = s, c["norm"], c["tol"]
    md.decomposed = 1 if c["decomposed"] else 0
    md.dd_type = 0
    md.dd_dim, md.dd_size = c["dd_dim"], c["dd_size"]
    md.decomposition = 0
    md.lossless = 2 if c.get("lossless") == "zstd" else 0
    md.huff_dict_size, md.huff_block_size = c["dict"], c["block"]
    md.backend = 3
    return md


def serialize(md):
    lib = _lib.load()
    n = ctypes.c_uint64()
    assert lib.mgx_metadata_serialize(ctypes.byref(md), None, 0, ctypes.byref(n)) == 2
    buf = np.zeros(n.value, dtype=np.uint8)
    assert lib.mgx_metadata_serialize(ctypes.byref(md), buf.ctypes.data, buf.size,
                                      ctypes.byref(n)) == 0
    return bytes(buf)


def parse(blob):
    lib = _lib.load()
    buf = np.frombuffer(blob, dtype=np.uint8)
    md = _lib.MgxMetadata()
    n = ctypes.c_uint64()
    rc = lib.mgx_metadata_parse(buf.ctypes.data, buf.size, ctypes.byref(md),
                                ctypes.byref(n))
    return rc, md, n.value


@pytest.mark.parametrize("name", sorted(GOLDEN))
def test_serialize_matches_protobuf_runtime(name):
    c = GOLDEN[name]
    assert serialize(make_md(c)).hex() == c["hex"]


@pytest.mark.parametrize("name", sorted(GOLDEN))
def test_parse_golden(name):
    c = GOLDEN[name]
    blob = bytes.fromhex(c["hex"])
    rc, md, n = parse(blob + b"\x00" * 32)   # records follow the header
    assert rc == 0 and n == len(blob)
    assert [md.shape[d] for d in range(md.D)] == c["shape"]
    assert md.dtype == (1 if c["dtype"] == "f64" else 0)
    assert md.mode == (0 if c["mode"] == "REL" else 1)
    s = math.inf if c["s"] == "inf" else float(c["s"])
    assert md.s == s and md.linf == (1 if math.isinf(s) else 0)
    assert md.tol == c["tol"]
    if c["mode"] == "REL":
        assert md.norm == c["norm"]
    assert bool(md.decomposed) == c["decomposed"]
    assert (md.dd_dim, md.dd_size) == (c["dd_dim"], c["dd_size"])
    assert (md.huff_dict_size, md.huff_block_size) == (c["dict"], c["block"])
    assert md.lossless == (2 if c.get("lossless") == "zstd" else 0)
    assert md.backend == 3 and md.decomposition == 0
    assert (md.version[0], md.file_version[0]) == (1, 0)


def test_crc32_is_zlib():
    lib = _lib.load()
    rng = np.random.default_rng(7)
    for n in (0, 1, 7, 64, 1000, 65537):
        b = rng.integers(0, 256, n, dtype=np.uint8)
        got = lib.mgx_crc32(b.ctypes.data if n else None, n)
        assert got == (zlib.crc32(bytes(b)) & 0xFFFFFFFF)


def reseal(blob, body):
    return (b"MGARD" + struct.pack("<Q", len(body)) +
            struct.pack("<I", zlib.crc32(body) & 0xFFFFFFFF) + body)


def test_parse_rejects_what_the_reference_rejects():
    blob = bytes.fromhex(GOLDEN["f32_3d_rel_linf"]["hex"])
    # signature (Metadata.cpp:101-110)
    assert parse(b"MGARX" + blob[5:])[0] != 0
    # CRC (Metadata.cpp:519-521)
    bad = bytearray(blob)
    bad[-1] ^= 1
    assert parse(bytes(bad))[0] != 0
    # truncated
    assert parse(blob[:-3])[0] != 0
    assert parse(blob[:10])[0] != 0
    body = blob[17:]
    # a newer major version (Metadata.cpp:531-533): mgard_version is field 2,
    # {major_ = 9}
    assert body[:4] == bytes([0x12, 0x02, 0x08, 0x01])
    assert parse(reseal(blob, bytes([0x12, 0x02, 0x08, 0x09]) + body[4:]))[0] != 0
    # CPU backend = "not compressed with MGARD-X" (Metadata.cpp:757-759): drop
    # the Device.backend field -> default 0
    assert body[-4:] == bytes([0x62, 0x02, 0x08, 0x03])
    assert parse(reseal(blob, body[:-4] + bytes([0x62, 0x00])))[0] != 0
    # unknown fields are skipped like protobuf does (field 15, varint)
    rc, md, _ = parse(reseal(blob, body + bytes([0x78, 0x05])))
    assert rc == 0 and md.D == 3


def test_unpacked_repeated_shape_is_accepted():
    # proto3 parsers must accept the non-packed encoding of repeated scalars
    c = GOLDEN["f64_2d_abs_l2"]
    blob = bytes.fromhex(c["hex"])
    body = blob[17:]
    packed = bytes([0x12, 0x04, 0x81, 0x01, 0x81, 0x02])     # shape = [129, 257]
    i = body.index(packed)
    unpacked = bytes([0x10, 0x81, 0x01, 0x10, 0x81, 0x02])
    grid_old = bytes([0x08, 0x02]) + packed
    grid_new = bytes([0x08, 0x02]) + unpacked
    j = body.index(grid_old)
    assert i == j + 2
    # Domain{ field 2: grid } : lengths stay the same (6 bytes either way)
    rc, md, _ = parse(reseal(blob, body.replace(grid_old, grid_new)))
    assert rc == 0 and [md.shape[0], md.shape[1]] == [129, 257]


def test_max_dim_by_memory_follows_the_halving_rule():
    """generate_max_dim_domain_decomposition_strategy, DomainDecomposer.hpp:192-223"""
    lib = _lib.load()
    shape = (ctypes.c_uint64 * 3)(1025, 1025, 1025)
    dec, dim = ctypes.c_int(), ctypes.c_int()
    size, num = ctypes.c_uint64(), ctypes.c_uint64()
    big = lib.mgx_estimate_footprint(3, shape, 0, math.inf, 0)
    assert big > 1025 ** 3 * 4
    args = (ctypes.byref(dec), ctypes.byref(dim), ctypes.byref(size), ctypes.byref(num))
    assert lib.mgx_max_dim_by_memory(3, shape, 0, math.inf, big + 1, *args) == 0
    assert (dec.value, dim.value, size.value, num.value) == (0, 0, 1025, 1)
    # does not fit: halve (round up) until one slab (with prefetch doubling) fits
    assert lib.mgx_max_dim_by_memory(3, shape, 0, math.inf, big, *args) == 0
    assert dec.value == 1 and dim.value == 0
    sz, want = 1025, None
    pre = 0
    while True:
        sub = (ctypes.c_uint64 * 3)(sz, 1025, 1025)
        if lib.mgx_estimate_footprint(3, sub, 0, math.inf, pre) < big:
            want = sz
            break
        sz = (sz - 1) // 2 + 1
        pre = 1 if (1025 - 1) // sz + 1 > 1 else 0
    assert size.value == want and num.value == (1025 - 1) // want + 1
    # largest dimension wins, first one on ties
    shape2 = (ctypes.c_uint64 * 3)(100, 400, 400)
    assert lib.mgx_max_dim_by_memory(3, shape2, 1, 0.0, 1 << 26, *args) == 0
    assert dec.value == 1 and dim.value == 1
    # nothing fits
    assert lib.mgx_max_dim_by_memory(3, shape2, 1, 0.0, 1000, *args) != 0


def test_config_defaults():
    from mgard_b200 import highlevel
    cfg = highlevel.Config()
    assert (cfg.huff_dict_size, cfg.huff_block_size) == (8192, 1024)
    assert cfg.dev_type == 0 and cfg.lossless == 0 and cfg.auto_pin_host_buffers == 1
    assert cfg.max_memory_footprint == 2 ** 64 - 1


def test_highlevel_refuses_without_a_device():
    import torch
    if torch.cuda.is_available():
        pytest.skip("this checks the GPU-less behaviour")
    from mgard_b200 import highlevel
    from mgard_b200.compressor import MgardError
    x = np.zeros((9, 9, 9), dtype=np.float32)
    with pytest.raises(MgardError) as e:
        highlevel.compress(x, 1e-3, math.inf, highlevel.REL)
    assert e.value.status == 5          # BackendNotAvailableFailure
    with pytest.raises(MgardError) as e:
        highlevel.compress(x, 1e-3, math.inf, highlevel.REL,
                           config=highlevel.Config(dev_type=highlevel.SERIAL))
    assert e.value.status == 5
