"""ctypes loaders for the two CPU oracles (TEST INFRASTRUCTURE ONLY).

* ``ref``  -- oracle/_ref/libmgardx_ref.so: the reference's own SERIAL/OPENMP
  MGARD-X low-level path compiled from /root/reference by oracle/Makefile.
* ``port`` -- oracle/libmgx_oracle.so: the plain-C restatement (mgx_oracle.c).

Nothing under mgard_b200/ imports this module.
"""
import ctypes
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
REF_SO = os.path.join(ORACLE_DIR, "_ref", "libmgardx_ref.so")
PORT_SO = os.path.join(ORACLE_DIR, "libmgx_oracle.so")

c_u64 = ctypes.c_uint64
c_dbl = ctypes.c_double
c_vp = ctypes.c_void_p


def _vp(a):
    return a.ctypes.data_as(c_vp)


def _shape(shape):
    return (c_u64 * len(shape))(*[int(s) for s in shape])


def _suffix(D, dtype):
    return "%d%s" % (D, "f" if np.dtype(dtype) == np.float32 else "d")


def _psuffix(dtype):
    return "f32" if np.dtype(dtype) == np.float32 else "f64"


def build_port():
    if not os.path.exists(PORT_SO) or os.path.getmtime(PORT_SO) < max(
        os.path.getmtime(os.path.join(ORACLE_DIR, f))
        for f in ("mgx_oracle.c", "mgx_oracle_core.inc",
                  "mgx_oracle_huffman.inc", "mgx_oracle.h")):
        subprocess.check_call(["make", "-C", ORACLE_DIR, "port"],
                              stdout=subprocess.DEVNULL)


_ref = None
_port = None


def have_ref():
    return os.path.exists(REF_SO)


def ref():
    global _ref
    if _ref is None:
        _ref = ctypes.CDLL(REF_SO)
    return _ref


def port():
    global _port
    if _port is None:
        build_port()
        _port = ctypes.CDLL(PORT_SO)
        for t in ("f32", "f64"):
            getattr(_port, "mgxo_norm_" + t).restype = c_dbl
        _port.mgxo_oob_reads.restype = c_u64
        _port.mgxo_set_oob_value.argtypes = [ctypes.c_uint32]
    return _port


import contextlib


@contextlib.contextmanager
def grid(coords=None, normalize=True, dtype=None):
    """Grid of every hierarchy the oracles build inside the block: ``coords`` = one
    1-D array of node coordinates per dimension (-> Hierarchy(shape, coords, config),
    a NON-UNIFORM grid; converted to ``dtype``, the data type of the calls), or None
    for a uniform grid with Config::normalize_coordinates = ``normalize``."""
    arrs, ptrs, D = None, None, 0
    if coords is not None:
        assert dtype is not None
        arrs = [np.ascontiguousarray(c, dtype) for c in coords]
        D = len(arrs)
        ptrs = (c_vp * D)(*[a.ctypes.data for a in arrs])

    def apply(p, D_, norm_, dt):
        port().mgxo_set_grid(D_, p, int(norm_))
        if have_ref():
            for d in range(1, 6):
                for t in ("f", "d"):
                    fn = getattr(ref(), "ref_set_grid_%d%s" % (d, t), None)
                    if fn is None:
                        continue
                    mine = p is not None and d == D_ and t == (
                        "f" if np.dtype(dt) == np.float32 else "d")
                    fn(p if mine else None, int(norm_))

    apply(ptrs, D, normalize, dtype)
    try:
        yield arrs
    finally:
        apply(None, 0, True, None)


class Oracle:
    """Uniform python face over either library."""

    def __init__(self, kind):
        assert kind in ("ref", "port")
        self.kind = kind
        self.lib = ref() if kind == "ref" else port()

    # -- stage: decomposition ------------------------------------------------
    def decompose(self, u, orthogonal):
        v = np.ascontiguousarray(u).copy()
        D = v.ndim
        if self.kind == "ref":
            fn = getattr(self.lib, "ref_decompose_" + _suffix(D, v.dtype))
            rc = fn(_shape(v.shape), _vp(v), int(orthogonal))
        else:
            fn = getattr(self.lib, "mgxo_decompose_" + _psuffix(v.dtype))
            rc = fn(D, _shape(v.shape), _vp(v), int(orthogonal))
        assert rc == 0
        return v

    def recompose(self, c, orthogonal):
        v = np.ascontiguousarray(c).copy()
        D = v.ndim
        if self.kind == "ref":
            fn = getattr(self.lib, "ref_recompose_" + _suffix(D, v.dtype))
            rc = fn(_shape(v.shape), _vp(v), int(orthogonal))
        else:
            fn = getattr(self.lib, "mgxo_recompose_" + _psuffix(v.dtype))
            rc = fn(D, _shape(v.shape), _vp(v), int(orthogonal))
        assert rc == 0
        return v

    # -- stage: quantization -------------------------------------------------
    def quantize(self, c, rel, tol, s, norm, dict_size=8192):
        c = np.ascontiguousarray(c)
        D = c.ndim
        codes = np.zeros(c.shape, np.int64)
        if self.kind == "ref":
            fn = getattr(self.lib, "ref_quantize_" + _suffix(D, c.dtype))
            rc = fn(_shape(c.shape), _vp(c), int(rel), c_dbl(tol), c_dbl(s),
                    c_dbl(norm), dict_size, _vp(codes))
        else:
            fn = getattr(self.lib, "mgxo_quantize_" + _psuffix(c.dtype))
            rc = fn(D, _shape(c.shape), _vp(c), int(rel), c_dbl(tol), c_dbl(s),
                    c_dbl(norm), dict_size, _vp(codes))
        assert rc == 0
        return codes

    def dequantize(self, codes, dtype, rel, tol, s, norm, dict_size=8192):
        codes = np.ascontiguousarray(codes, np.int64)
        D = codes.ndim
        c = np.zeros(codes.shape, dtype)
        if self.kind == "ref":
            fn = getattr(self.lib, "ref_dequantize_" + _suffix(D, dtype))
            rc = fn(_shape(c.shape), _vp(codes), int(rel), c_dbl(tol),
                    c_dbl(s), c_dbl(norm), dict_size, _vp(c))
        else:
            fn = getattr(self.lib, "mgxo_dequantize_" + _psuffix(dtype))
            rc = fn(D, _shape(c.shape), _vp(codes), int(rel), c_dbl(tol),
                    c_dbl(s), c_dbl(norm), dict_size, _vp(c))
        assert rc == 0
        return c

    # -- stage: Huffman ------------------------------------------------------
    def huffman_compress(self, codes, dict_size=8192, chunk=1024):
        codes = np.ascontiguousarray(codes, np.int64).ravel()
        n = codes.size
        cap = 64 + 16 * (n // chunk + 1) + 8 * (128 + dict_size) + 24 * n + 4096
        out = np.zeros(cap, np.uint8)
        nb = c_u64(0)
        name = ("ref_huffman_compress" if self.kind == "ref" else
                "mgxo_huffman_compress")
        rc = getattr(self.lib, name)(_vp(codes), c_u64(n), dict_size, chunk,
                                     _vp(out), c_u64(cap), ctypes.byref(nb))
        assert rc == 0, rc
        return out[:nb.value].copy()

    def huffman_decompress(self, blob, n, dict_size=8192, chunk=1024):
        blob = np.ascontiguousarray(blob, np.uint8)
        codes = np.zeros(n, np.int64)
        if self.kind == "ref":
            rc = self.lib.ref_huffman_decompress(_vp(blob), c_u64(blob.size),
                                                 c_u64(n), dict_size, chunk,
                                                 _vp(codes))
        else:
            rc = self.lib.mgxo_huffman_decompress(_vp(blob), c_u64(blob.size),
                                                  c_u64(n), _vp(codes))
        assert rc == 0, rc
        return codes

    # -- whole path ----------------------------------------------------------
    def compress(self, u, rel, tol, s, norm=0.0, dict_size=8192, chunk=1024,
                 backend=0):
        u = np.ascontiguousarray(u)
        D = u.ndim
        cap = u.nbytes * 3 + (1 << 20)
        out = np.zeros(cap, np.uint8)
        nb = c_u64(0)
        nrm = c_dbl(norm)
        if self.kind == "ref":
            fn = getattr(self.lib, "ref_compress_" + _suffix(D, u.dtype))
            rc = fn(backend, _shape(u.shape), _vp(u), int(rel), c_dbl(tol),
                    c_dbl(s), ctypes.byref(nrm), dict_size, chunk, _vp(out),
                    c_u64(cap), ctypes.byref(nb))
        else:
            fn = getattr(self.lib, "mgxo_compress_" + _psuffix(u.dtype))
            rc = fn(D, _shape(u.shape), _vp(u), int(rel), c_dbl(tol), c_dbl(s),
                    ctypes.byref(nrm), dict_size, chunk, _vp(out), c_u64(cap),
                    ctypes.byref(nb))
        assert rc == 0, rc
        return out[:nb.value].copy(), nrm.value

    def decompress(self, blob, shape, dtype, rel, tol, s, norm,
                   dict_size=8192, chunk=1024, backend=0):
        blob = np.ascontiguousarray(blob, np.uint8)
        out = np.zeros(shape, dtype)
        D = out.ndim
        if self.kind == "ref":
            fn = getattr(self.lib, "ref_decompress_" + _suffix(D, dtype))
            rc = fn(backend, _shape(shape), _vp(blob), c_u64(blob.size),
                    int(rel), c_dbl(tol), c_dbl(s), c_dbl(norm), dict_size,
                    chunk, _vp(out))
        else:
            fn = getattr(self.lib, "mgxo_decompress_" + _psuffix(dtype))
            rc = fn(D, _shape(shape), _vp(blob), c_u64(blob.size), int(rel),
                    c_dbl(tol), c_dbl(s), c_dbl(norm), dict_size, _vp(out))
        assert rc == 0, rc
        return out


def smooth_field(shape, dtype, seed=42, noise=1e-6):
    """The reference's own test generator (tests/mgard-x/src/
    test_Compressor.cpp.in:27-43) plus a higher-frequency term and a small
    integer-hash noise -- closed form, libm-only."""
    idx = np.indices(shape, dtype=np.float64)
    D = len(shape)
    v = np.ones(shape, np.float64)
    ph = np.zeros(shape, np.float64)
    for d in range(D):
        v *= np.sin(2.0 * np.pi * idx[d] / shape[d] + 0.5 * d)
        ph += (d + 1) * idx[d] / shape[d]
    for k in range(1, 5):
        v += 0.05 * 2.0 ** (-k) * np.sin(2.0 * np.pi * (2 * k + 1) * ph)
    lin = np.arange(int(np.prod(shape)), dtype=np.uint64).reshape(shape)
    h = (lin + np.uint64(seed)) * np.uint64(0x9E3779B97F4A7C15)
    h ^= h >> np.uint64(31)
    h *= np.uint64(0xBF58476D1CE4E5B9)
    h ^= h >> np.uint64(29)
    v += noise * ((h >> np.uint64(11)).astype(np.float64) / 2.0 ** 53 - 0.5)
    return v.astype(dtype)


def fresh(kind, method, *args, **kwargs):
    """Run Oracle(kind).method(*args) in a NEW process.

    The reference's GenerateCL reads one element past its frequency array
    (see oracle/mgx_oracle_huffman.inc); what it finds there depends on the
    heap's history, so bit-exact comparisons against the compiled reference
    are only reproducible from a fresh process (fresh heap => 0)."""
    import pickle
    import sys
    import tempfile
    with tempfile.TemporaryDirectory() as td:
        pin, pout = os.path.join(td, "in.pkl"), os.path.join(td, "out.pkl")
        with open(pin, "wb") as f:
            pickle.dump((kind, method, args, kwargs), f)
        code = ("import pickle,sys; sys.path.insert(0,%r); import oracle_lib as o;"
                "k,m,a,kw=pickle.load(open(%r,'rb'));"
                "r=getattr(o.Oracle(k),m)(*a,**kw);"
                "pickle.dump(r,open(%r,'wb'))" % (os.path.dirname(__file__), pin, pout))
        subprocess.check_call([sys.executable, "-c", code])
        with open(pout, "rb") as f:
            return pickle.load(f)


def fresh_in_grid(kind, coords, dtype, normalize, method, *args, **kwargs):
    """fresh() inside ``grid(coords, normalize, dtype)``"""
    import pickle
    import sys
    import tempfile
    with tempfile.TemporaryDirectory() as td:
        pin, pout = os.path.join(td, "in.pkl"), os.path.join(td, "out.pkl")
        with open(pin, "wb") as f:
            pickle.dump((kind, coords, dtype, normalize, method, args, kwargs), f)
        code = ("import pickle,sys; sys.path.insert(0,%r); import oracle_lib as o;"
                "k,c,dt,nz,m,a,kw=pickle.load(open(%r,'rb'));\n"
                "with o.grid(c, nz, dt):\n"
                "    r=getattr(o.Oracle(k),m)(*a,**kw)\n"
                "pickle.dump(r,open(%r,'wb'))" % (os.path.dirname(__file__), pin, pout))
        subprocess.check_call([sys.executable, "-c", code])
        with open(pout, "rb") as f:
            return pickle.load(f)


def same_container(a, b):
    """byte equality except byte 7 (uninitialised pad after the 7-byte
    signature in the reference's Serialize)."""
    return a.size == b.size and np.array_equal(np.delete(a, 7), np.delete(b, 7))


def ref_allowed_outputs(method, *args, **kwargs):
    """Outputs of the restatement for the outcomes the reference's
    out-of-bounds read allows (stand-in value 0 = the defined behaviour, and
    UINT32_MAX = 'garbage larger than any frequency'); one element if the
    read is never hit for this input."""
    p = Oracle("port")
    p.lib.mgxo_set_oob_value(0)
    p.lib.mgxo_oob_reads()
    outs = [getattr(p, method)(*args, **kwargs)]
    if p.lib.mgxo_oob_reads() > 0:
        p.lib.mgxo_set_oob_value(0xFFFFFFFF)
        try:
            outs.append(getattr(p, method)(*args, **kwargs))
        finally:
            p.lib.mgxo_set_oob_value(0)
    return outs
