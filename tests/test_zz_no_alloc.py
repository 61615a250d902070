"""mgx_norm_partial_noalloc (no counterpart in the reference):
the allocation-free norm reduction, called
through the C ABI, against the existing mgx_norm_partial on the same inputs --
two calls in a row on the shared static scratch, the second one shorter (stale
scratch entries of the first call must not leak in) -- and its size limit.

Collected last on purpose (file name)."""
import ctypes

import numpy as np
import pytest
import torch

import mgard_b200 as mg
from mgard_b200 import _lib

pytestmark = pytest.mark.gpu

MGX_OUTPUT_TOO_LARGE = 2


def _noalloc(lib, t, linf, out):
    err = ctypes.create_string_buffer(256)
    stream = torch.cuda.current_stream(t.device)
    rc = lib.mgx_norm_partial_noalloc(
        t.device.index or 0, ctypes.c_void_p(stream.cuda_stream), t.data_ptr(),
        0 if t.dtype == torch.float32 else 1, t.numel(), int(linf), out.data_ptr(),
        err, 256)
    assert rc == 0, err.value.decode()
    stream.synchronize()
    return float(out.item())


@pytest.mark.parametrize("tdt", [torch.float32, torch.float64])
@pytest.mark.parametrize("linf", [1, 0])
def test_noalloc_norm_matches_the_context_entry_point_twice(tdt, linf):
    lib = _lib.load()
    dev = torch.device("cuda", 0)
    out = torch.zeros(1, dtype=torch.float64, device=dev)
    g = torch.Generator(device="cpu").manual_seed(5)
    # first call: 5.3 M elements (6 scratch entries); second: 70 k (1 entry)
    for shape in ((129, 200, 205), (70, 33, 31)):
        u = (torch.randn(shape, generator=g, dtype=torch.float64) * 3.0 + 0.25).to(tdt).to(dev)
        comp = mg.Compressor(mg.Hierarchy(shape, tdt))
        want = comp.NormPartial(u, float("inf") if linf else 0.0)
        out.fill_(-1.0)
        torch.cuda.synchronize()
        got = _noalloc(lib, u, linf, out)
        if linf:
            assert got == want == float(u.abs().max().item())
        else:
            # the tolerance the norm tests use (test_gpu_parity: 1e-6 relative);
            # the two entry points only differ in the order of fp64 partial sums
            assert abs(got - want) <= 1e-6 * abs(want), (got, want)
            assert abs(got - want) <= 1e-12 * abs(want), (got, want)
            # and against a sum that shares no code with the library
            indep = float((u.double() ** 2).sum().item())
            assert abs(got - indep) <= 1e-12 * indep, (got, indep)
        # repeated call, same input: bit-identical (fixed summation order)
        assert _noalloc(lib, u, linf, out) == got


def test_noalloc_norm_rejects_sizes_beyond_the_scratch():
    lib = _lib.load()
    out = torch.zeros(1, dtype=torch.float64, device="cuda")
    small = torch.ones(16, dtype=torch.float32, device="cuda")
    err = ctypes.create_string_buffer(256)
    rc = lib.mgx_norm_partial_noalloc(0, None, small.data_ptr(), 0, (1 << 31) + 1, 1,
                                      out.data_ptr(), err, 256)
    assert rc == MGX_OUTPUT_TOO_LARGE
    assert b"exceed" in err.value and b"2147483648" in err.value
    # the limit itself is not rejected for its size (bad dtype is reported first)
    rc = lib.mgx_norm_partial_noalloc(0, None, small.data_ptr(), 7, 1 << 31, 1,
                                      out.data_ptr(), err, 256)
    assert rc == 4 and b"dtype" in err.value


def test_noalloc_norm_replays_from_a_cuda_graph():
    """captured once (after one eager call on the device), replayed twice on changing
    input: what a caller under CUDA graph capture does"""
    lib = _lib.load()
    dev = torch.device("cuda", 0)
    g = torch.Generator(device="cpu").manual_seed(9)
    u = (torch.randn((64, 70, 130), generator=g, dtype=torch.float64) + 0.5).to(torch.float32).to(dev)
    out = torch.zeros(1, dtype=torch.float64, device=dev)
    eager = _noalloc(lib, u, 0, out)  # resolves the scratch address, loads the module
    s = torch.cuda.Stream(device=dev)
    graph = torch.cuda.CUDAGraph()
    err = ctypes.create_string_buffer(256)
    s.wait_stream(torch.cuda.current_stream(dev))
    with torch.cuda.graph(graph, stream=s):
        rc = lib.mgx_norm_partial_noalloc(
            0, ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream), u.data_ptr(), 0,
            u.numel(), 0, out.data_ptr(), err, 256)
    assert rc == 0, err.value.decode()
    out.fill_(-1.0)
    graph.replay()
    torch.cuda.synchronize()
    assert float(out.item()) == eager
    u.mul_(2.0)  # exact in floating point: the sum of squares is exactly 4 x
    graph.replay()
    torch.cuda.synchronize()
    assert float(out.item()) == 4.0 * eager
