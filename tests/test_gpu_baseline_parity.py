"""Parity at the sizes BASELINE.md ran on the oracle, and the robustness of the decoder.

* cfg 1 (513^3 fp32 rel L-inf 1e-3): container bytes == the reference's own SERIAL
  build (oracle/_ref, fresh process), and the reference decodes our stream;
* 1025^2 fp64 L2 and 8x16x20x20x101 fp64 s-norm: bytes == oracle with the oracle's norm;
* REL + L2 / s-norm END TO END WITH THE GPU'S OWN NORM: codes against the oracle run
  with the oracle's norm -- mismatches counted, each off by exactly one code (a
  rounding-boundary tie moved by the last bits of the norm), CR within 1 %;
* a container with its outlier list in append order (what the reference's CUDA / OpenMP
  back ends write) decodes to the same field; corrupt chunk offsets, outlier indices
  and decodebook keys are refused instead of read out of bounds.
"""
import numpy as np
import pytest
import torch

import oracle_lib
from oracle_lib import Oracle, smooth_field, fresh, same_container

import mgard_b200 as mg

pytestmark = pytest.mark.gpu
INF = float("inf")


def make(shape, dtype):
    tdt = torch.float32 if dtype == np.float32 else torch.float64
    return mg.Compressor(mg.Hierarchy(shape, tdt))


def test_cfg1_513cubed_bytes_equal_the_reference_serial_build():
    shape, tol = (513, 513, 513), 1e-3
    u = smooth_field(shape, np.float32)
    kind = "ref" if oracle_lib.have_ref() else "port"
    want, wnorm = fresh(kind, "compress", u, 1, tol, INF)  # backend 0 = SERIAL
    comp = make(shape, np.float32)
    blob, norm = comp.Compress(torch.from_numpy(u).cuda(), mg.REL, tol, INF)
    got = blob.cpu().numpy()
    assert norm == wnorm
    assert same_container(got, want), "513^3 container differs from " + kind
    # and the reference decodes our bytes within the bound
    back = fresh(kind, "decompress", got, shape, np.float32, 1, tol, INF, norm)
    assert np.abs(back.astype(np.float64) - u).max() <= tol * norm


@pytest.mark.parametrize("shape,dtype,s,tol", [((1025, 1025), np.float64, 0.0, 1e-4),
                                               ((8, 16, 20, 20, 101), np.float64, 1.0, 1e-2),
                                               ((513, 513), np.float32, 0.0, 1e-4)])
def test_baseline_probe_sizes_bytes_equal_the_oracle(shape, dtype, s, tol):
    u = smooth_field(shape, dtype)
    want, wnorm = Oracle("port").compress(u, 1, tol, s)
    comp = make(shape, dtype)
    blob, _ = comp.Compress(torch.from_numpy(u).cuda(), mg.REL, tol, s, norm=wnorm)
    assert same_container(blob.cpu().numpy(), want)


@pytest.mark.parametrize("shape,dtype,s,tol", [((1025, 1025), np.float64, 0.0, 1e-4),
                                               ((513, 513), np.float32, 0.0, 1e-4),
                                               ((65, 70, 130), np.float32, 1.0, 1e-3),
                                               ((6, 10, 12, 14, 40), np.float64, 1.0, 1e-2)])
def test_rel_l2_end_to_end_with_the_gpus_own_norm(shape, dtype, s, tol):
    """north_star: code mismatches confined to rounding-boundary ties and counted,
    compression ratio within 1 % of the reference."""
    u = smooth_field(shape, dtype)
    o = Oracle("port")
    want_blob, wnorm = o.compress(u, 1, tol, s)
    want_codes = o.quantize(o.decompose(u, 1), 1, tol, s, wnorm)
    comp = make(shape, dtype)
    d = torch.from_numpy(u).cuda()
    blob, gnorm = comp.Compress(d, mg.REL, tol, s)  # norm computed on the GPU
    rel_norm = abs(gnorm - wnorm) / wnorm
    # fp64: both sums are fp64, only their order differs.  fp32: the reference sums
    # the squares in float (NormCalculator.hpp:35-64, a different value on each of its
    # back ends: a serial float sum of 6e5 squares is already 3e-4 off); the GPU
    # accumulates in fp64
    assert rel_norm <= (1e-12 if dtype == np.float64 else 1e-3)
    coeff = d.clone()
    comp.Decompose(coeff, 1)
    got_codes = comp.Quantize(coeff, mg.REL, tol, s, gnorm).cpu().numpy()
    diff = got_codes - want_codes
    bad = np.flatnonzero(diff)
    # a relative norm difference r moves |t| = |c * quantizer| by r * |t|: a code can
    # change by one when |t| sits within that of a half-integer (and by about r * |t|
    # where that exceeds one: the huge codes of the coarsest levels, stored as outliers)
    mag = np.abs(want_codes.astype(np.float64) - 4096.0)
    assert np.all(np.abs(diff.ravel()[bad]) <= 1 + 2 * rel_norm * mag.ravel()[bad])
    # P(code changes) ~ 2 r |t| for a uniformly distributed fractional part
    expected = 2.0 * rel_norm * float(np.minimum(mag, 1e6).sum())
    assert bad.size <= 10 + 3 * expected, (bad.size, expected, rel_norm)
    cr_ref = u.nbytes / want_blob.size
    cr = u.nbytes / blob.numel()
    assert abs(cr - cr_ref) <= 0.01 * cr_ref
    # error bound on the reconstruction
    dec = comp.Decompress(blob, mg.REL, tol, s, gnorm).cpu().numpy().astype(np.float64)
    err = np.sqrt(((dec - u.astype(np.float64)) ** 2).mean())
    assert err <= tol * gnorm
    print("own-norm run %s: norm rel. diff %.2e, %d of %d codes differ (ties)"
          % (shape, rel_norm, bad.size, u.size))


# ---- decoder robustness ------------------------------------------------------
def _sections(blob, n, dict_size=8192, chunk=1024):
    nchunk = (n - 1) // chunk + 1
    off_dbs = 32 + 16 * nchunk
    off_db = off_dbs + 8
    off_ddata = off_db + 8 * (128 + dict_size)
    ddata = int(blob[off_ddata:off_ddata + 8].view(np.uint64)[0])
    off_outc = off_ddata + 8 + 8 * ddata
    n_out = int(blob[off_outc:off_outc + 8].view(np.uint64)[0])
    return dict(nchunk=nchunk, off_bits=32, off_offs=32 + 8 * nchunk, off_db=off_db,
                off_keys=off_db + 8 * 128, ddata=ddata, off_outc=off_outc, n_out=n_out,
                off_oidx=off_outc + 8, off_oval=off_outc + 8 + 8 * n_out)


def _outlier_case():
    shape, tol = (65, 65, 129), 3e-6  # tight bound: thousands of outliers
    u = smooth_field(shape, np.float32)
    comp = make(shape, np.float32)
    out = torch.empty(6 * u.nbytes + (1 << 20), dtype=torch.uint8, device="cuda")
    blob, norm = comp.Compress(torch.from_numpy(u).cuda(), mg.REL, tol, INF, out=out)
    return shape, tol, u, comp, blob.cpu().numpy().copy(), norm


def test_unsorted_outlier_list_decodes_to_the_same_field():
    shape, tol, u, comp, blob, norm = _outlier_case()
    sec = _sections(blob, u.size)
    assert sec["n_out"] > 100
    want = comp.Decompress(torch.from_numpy(blob).cuda(), mg.REL, tol, INF, norm).cpu().numpy()
    perm = np.random.default_rng(3).permutation(sec["n_out"])
    idx = blob[sec["off_oidx"]:sec["off_oval"]].view(np.uint64).copy()
    val = blob[sec["off_oval"]:sec["off_oval"] + 8 * sec["n_out"]].view(np.int64).copy()
    shuffled = blob.copy()
    shuffled[sec["off_oidx"]:sec["off_oval"]] = idx[perm].view(np.uint8)
    shuffled[sec["off_oval"]:sec["off_oval"] + 8 * sec["n_out"]] = val[perm].view(np.uint8)
    got = comp.Decompress(torch.from_numpy(shuffled).cuda(), mg.REL, tol, INF, norm).cpu().numpy()
    assert np.array_equal(got, want)
    assert np.abs(got.astype(np.float64) - u).max() <= tol * norm


@pytest.mark.parametrize("what", ["chunk_offset", "chunk_bits", "outlier_index", "key"])
def test_corrupt_payload_is_refused(what):
    shape, tol, u, comp, blob, norm = _outlier_case()
    sec = _sections(blob, u.size)
    bad = blob.copy()
    if what == "chunk_offset":
        bad[sec["off_offs"] + 8 * 5:sec["off_offs"] + 8 * 6] = \
            np.array([sec["ddata"] + 7], np.uint64).view(np.uint8)
    elif what == "chunk_bits":
        bad[sec["off_bits"] + 8 * 3:sec["off_bits"] + 8 * 4] = \
            np.array([1 << 40], np.uint64).view(np.uint8)
    elif what == "outlier_index":
        bad[sec["off_oidx"] + 8 * 2:sec["off_oidx"] + 8 * 3] = \
            np.array([u.size + 12345], np.uint64).view(np.uint8)
    else:
        bad[sec["off_keys"] + 8 * 9:sec["off_keys"] + 8 * 10] = \
            np.array([70000], np.uint64).view(np.uint8)
    with pytest.raises(mg.MgardError):
        comp.Decompress(torch.from_numpy(bad).cuda(), mg.REL, tol, INF, norm)
    # the context is still usable afterwards
    ok = comp.Decompress(torch.from_numpy(blob).cuda(), mg.REL, tol, INF, norm).cpu().numpy()
    assert np.abs(ok.astype(np.float64) - u).max() <= tol * norm
