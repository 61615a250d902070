"""Partitioned Thomas solves of the correction path (mgx_level.cu: k_ipk_part_contig /
k_ipk_part_strided; replace the one-thread-per-chain Ipk1/2/3Reo sweeps,
IterativeProcessingKernel3D.hpp:28-1169, IPKFunctor.h:13-50) on chains of >= 1024
elements:

* containers byte-identical to the oracle's (which solves every chain sequentially),
* byte-identical to the sequential GPU kernels (MGX_NO_PART=1) at sizes the oracle
  does not reach, with no chain redone,
* a field whose dynamic range defeats the warm-up (1e+250 next to 1e-250) is caught by
  the in-kernel check, redone sequentially, and still byte-identical.
"""
import numpy as np
import pytest
import torch

from oracle_lib import Oracle, smooth_field, same_container

import mgard_b200 as mg

pytestmark = pytest.mark.gpu


def make(shape, dtype):
    tdt = torch.float32 if dtype == np.float32 else torch.float64
    return mg.Compressor(mg.Hierarchy(shape, tdt))


def big_out(u):
    return torch.empty(3 * u.nbytes + (1 << 20), dtype=torch.uint8, device="cuda")


@pytest.mark.parametrize("shape,dtype,s,tol", [
    ((1025, 2100), np.float64, 0.0, 1e-4),     # strided and contiguous solves, ragged
    ((2049, 1030), np.float32, 1.0, 1e-3),     # only the strided one is long enough
    ((5, 9, 4100), np.float64, 0.0, 1e-3),     # 3-D, one long contiguous axis
    ((2500, 6, 7), np.float32, 0.0, 1e-3),     # 3-D, long strided axis with inner = 42
])
def test_containers_equal_the_oracle(shape, dtype, s, tol):
    u = smooth_field(shape, dtype)
    want, wnorm = Oracle("port").compress(u, 1, tol, s)
    comp = make(shape, dtype)
    blob, _ = comp.Compress(torch.from_numpy(u).cuda(), mg.REL, tol, s, norm=wnorm,
                            out=big_out(u))
    assert same_container(blob.cpu().numpy(), want)
    assert comp.LastRedoCount() == 0
    dec = comp.Decompress(blob, mg.REL, tol, s, wnorm).cpu().numpy().astype(np.float64)
    err = np.sqrt(((dec - u.astype(np.float64)) ** 2).mean())
    assert err <= tol * wnorm


@pytest.mark.parametrize("shape,dtype", [((4097, 4097), np.float64),
                                         ((3000, 5000), np.float32),
                                         ((6, 40, 40, 2300), np.float64)])
def test_bytes_equal_the_sequential_kernels(shape, dtype, monkeypatch):
    rng = np.random.default_rng(1)
    u = (smooth_field(shape, np.float64) + 1e-3 * rng.standard_normal(shape)).astype(dtype)
    d = torch.from_numpy(u).cuda()
    comp = make(shape, dtype)
    blob, norm = comp.Compress(d, mg.REL, 1e-3, 0.0, out=big_out(u))
    assert comp.LastRedoCount() == 0
    monkeypatch.setenv("MGX_NO_PART", "1")
    seq = make(shape, dtype)
    monkeypatch.delenv("MGX_NO_PART")
    blob_seq, norm_seq = seq.Compress(d, mg.REL, 1e-3, 0.0, out=big_out(u))
    assert norm == norm_seq
    assert torch.equal(blob, blob_seq)
    # decompress: partitioned against sequential, bit for bit
    a = comp.Decompress(blob, mg.REL, 1e-3, 0.0, norm)
    b = seq.Decompress(blob, mg.REL, 1e-3, 0.0, norm)
    assert torch.equal(a, b)


def test_huge_dynamic_range_is_caught_and_redone(monkeypatch):
    """a chain that holds 1e+250 a few elements before 1e-250: the warm-up of the
    partition after the jump starts from zero and cannot know the 1e+250 tail"""
    shape = (1500, 2100)
    u = smooth_field(shape, np.float64) * 1e-250
    u[700:, :1000] = 1e250   # jumps along both axes
    u[:, 1990:2040] *= 1e200
    d = torch.from_numpy(u).cuda()
    comp = make(shape, np.float64)
    coeff = d.clone()
    comp.Decompose(coeff, 1)
    monkeypatch.setenv("MGX_NO_PART", "1")
    seq = make(shape, np.float64)
    monkeypatch.delenv("MGX_NO_PART")
    coeff_seq = d.clone()
    seq.Decompose(coeff_seq, 1)
    assert torch.equal(coeff.view(torch.int64), coeff_seq.view(torch.int64))
    assert comp.LastRedoCount() > 0, "the field was meant to defeat the warm-up"
    back = coeff.clone()
    comp.Recompose(back, 1)
    back_seq = coeff_seq.clone()
    seq.Recompose(back_seq, 1)
    assert torch.equal(back.view(torch.int64), back_seq.view(torch.int64))
