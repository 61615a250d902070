"""mgx_x_compress_multi: one process, one host thread per GPU, MaxDim subdomains dealt
to the devices.  The stream must be byte-identical to what mgx_x_compress writes for
the same decomposition (GPUPipelines.hpp:89-205, Metadata.cpp:353-372) and decode on
one GPU.  The two-device cases are skipped on a one-GPU box."""
import ctypes

import numpy as np
import pytest
import torch

from oracle_lib import smooth_field

from mgard_b200 import _lib, highlevel as hl

pytestmark = pytest.mark.gpu
INF = float("inf")


def _single_call_with_slabs(u, tol, s, mode, rows, dim):
    """mgx_x_compress with a memory budget that makes the halving rule stop at `rows`"""
    lib = _lib.load()
    shp = list(u.shape)
    shp[dim] = rows
    code = 0 if u.dtype == np.float32 else 1
    fit = lib.mgx_estimate_footprint(u.ndim, (ctypes.c_uint64 * u.ndim)(*shp), code,
                                     ctypes.c_double(s), 1) + 1
    return hl.compress(u, tol, s, mode, config=hl.Config(max_memory_footprint=int(fit)))


@pytest.mark.parametrize("shape,dtype,s,tol,mode", [
    ((33, 40, 65), np.float32, INF, 1e-3, hl.REL),
    ((40, 50), np.float64, 0.0, 1e-4, hl.REL)])
def test_one_device_equals_the_plain_call(shape, dtype, s, tol, mode):
    u = smooth_field(shape, dtype, seed=2)
    a = hl.compress_multi(u, tol, s, mode, [0])
    b = np.asarray(hl.compress(u, tol, s, mode))
    assert a.size == b.size and np.array_equal(a, b)
    assert np.array_equal(hl.decompress_multi(a, [0]), hl.decompress(b))


def test_one_device_decodes_a_decomposed_stream_like_the_plain_call():
    u = smooth_field((70, 20, 24), np.float32, seed=4)
    stream = np.asarray(_single_call_with_slabs(u, 1e-3, INF, hl.REL, 18, 0))
    md, _ = hl.read_metadata(bytes(stream[:4096]))
    assert md.decomposed
    assert np.array_equal(hl.decompress_multi(stream, [0]), hl.decompress(stream))


@pytest.mark.parametrize("shape,dtype,s,tol,mode", [
    ((130, 72, 80), np.float32, INF, 1e-3, hl.REL),   # REL L-inf: MAX over the devices
    ((130, 33, 40), np.float64, 0.0, 1e-3, hl.REL),   # REL L2: SUM, sqrt(tol^2 / 2)
    ((6, 8, 9, 9, 200), np.float64, 1.0, 1e-2, hl.ABS),  # slabs along the last axis
])
def test_two_devices_equal_the_single_call_stream(shape, dtype, s, tol, mode):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    u = smooth_field(shape, dtype, seed=5)
    u[tuple(0 for _ in shape)] = 5.0  # the extreme value lives in device 0's block
    multi = hl.compress_multi(u, tol, s, mode, [0, 1])
    md, _ = hl.read_metadata(bytes(multi[:4096]))
    assert md.decomposed and md.dd_size == (max(shape) - 1) // 2 + 1
    single = np.asarray(_single_call_with_slabs(u, tol, s, mode, int(md.dd_size),
                                                int(md.dd_dim)))
    assert multi.size == single.size and np.array_equal(multi, single)
    back = hl.decompress(multi)
    # the records decoded on two devices, one host thread each: the same array
    assert np.array_equal(hl.decompress_multi(multi, [0, 1]), back)
    assert np.array_equal(hl.decompress_multi(multi, [1]), back)
    err = back.astype(np.float64) - u
    bound = tol * (md.norm if mode == hl.REL else 1.0)
    assert (np.abs(err).max() if np.isinf(s) else np.sqrt((err ** 2).mean())) <= bound


def test_a_device_listed_twice_is_refused():
    u = smooth_field((20, 20), np.float32)
    with pytest.raises(Exception):
        hl.compress_multi(u, 1e-3, INF, hl.REL, [0, 0])
