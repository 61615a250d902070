#!/usr/bin/env python
"""bench.py -- MGARD-X hot path (decompose + quantize + Huffman) on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
                    [--config auto|cfg3|cfg5|cfg4] [--tol T]

Workloads (BASELINE.json ``configs``; SURVEY.md section 8d generators):

* ``cfg3`` (default at N = 1): 3-D fp32 1025^3 NYX-shaped field (rho = exp(2.5 g), g = 32
  plane waves from splitmix64(42)), rel L_inf 1e-3.  One "step" = one Compress() of it.
  With N > 1 every rank holds its own cube (weak scaling, one scalar all-reduce).
* ``cfg5`` (default at N > 1): 3-D fp32 2049^3 smooth field, rel L_inf 1e-4, MGARD-X's
  MaxDim decomposition (DomainDecomposer.hpp:118-133, 192-223): seven 257-row slabs + one
  of 250 rows, shared out over the ranks in contiguous blocks (8/N slabs each: STRONG
  scaling, 34.4 GB whatever N is), one NCCL all-reduce(MAX) of the value range, every
  slab then compressed with the ABS bound tol * range (calc_local_abs_tol).
* ``cfg4``: 5-D fp64 8x16x39x39x39969 XGC-shaped field, rel s-norm (s = 1), split along
  the last axis into 8 slabs (7 x 4997 + 4990), all-reduce(SUM) of sum u^2.

Prints ONE JSON line (rank 0):

* ``value``     whole-job compress GB/s of INPUT bytes, inputs resident in HBM;
* ``e2e``       same metric through the host-pointer call (mgard_x::compress ->
                mgx_x_compress at N = 1; mgx_compress_host per slab otherwise): pinned host
                input -> H2D -> compress -> D2H of the compressed bytes inside the timed
                region;
* ``roofline``  ``step_frac`` = (N sizeof(T) + C) / t / peak, SURVEY 8(d)'s strict figure;
                ``frac`` = the dominant stage's algorithmic bytes / its CUDA-event time /
                peak; ``pass_model_frac`` = bytes of this implementation's passes / t / peak;
* ``cpu_baseline`` / ``--impl reference``: the reference's own OPENMP Compressor
                (oracle/_ref, built by oracle/Makefile from /root/reference) on the box's
                host cores, on the whole field when that fits the time budget, else on a
                stated corner of it; falls back to the C restatement (kind "port", 1
                core).  These are the only places this file touches oracle/.
* ``other_configs`` (N = 1): short runs of the remaining configurations, each with its own
                strict roofline fraction, dominant stage and CPU baseline.
"""
import argparse
import ctypes
import json
import math
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "compress GB/s (input bytes)"
UNIT = "GB/s"
L2_BYTES = 126 << 20
INF = float("inf")


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", default="auto", choices=["auto", "cfg3", "cfg5", "cfg4"])
    ap.add_argument("--edge", type=int, default=1025, help="cfg3: cube edge")
    ap.add_argument("--field", default="nyx", choices=["nyx", "smooth"],
                    help="cfg3: generator")
    ap.add_argument("--tol", type=float, default=None,
                    help="relative bound (default: 1e-3 cfg3, 1e-4 cfg5, 1e-3 cfg4)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-numa", action="store_true",
                    help="N > 1: do not bind each rank to its GPU's NUMA node")
    ap.add_argument("--no-stream-check", action="store_true",
                    help="N > 1: skip gathering the ranks' records into one MGARD stream and "
                         "decoding it on rank 0")
    ap.add_argument("--no-extra", action="store_true", help="skip other_configs")
    return ap.parse_args()


# ---------------------------------------------------------------------------
# synthetic fields: closed form + integer-hash noise, any index box of the global
# grid, on any torch device
# ---------------------------------------------------------------------------
def _splitmix64(state):
    state = (state + 0x9E3779B97F4A7C15) & 0xFFFFFFFFFFFFFFFF
    z = state
    z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & 0xFFFFFFFFFFFFFFFF
    z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & 0xFFFFFFFFFFFFFFFF
    return state, z ^ (z >> 31)


def _nyx_waves(seed=42, nwaves=32):
    """integer wave vectors in [-12, 12]^3 \\ {0}, phases, amplitudes ~ |k|^-1.5 scaled to
    unit variance of g"""
    st = seed
    ks, ph = [], []
    while len(ks) < nwaves:
        k = []
        for _ in range(3):
            st, r = _splitmix64(st)
            k.append(int(r % 25) - 12)
        if k == [0, 0, 0]:
            continue
        st, r = _splitmix64(st)
        ks.append(k)
        ph.append(2.0 * math.pi * (r >> 11) / 2.0 ** 53)
    ks = np.array(ks, np.float64)
    amp = np.linalg.norm(ks, axis=1) ** -1.5
    amp *= math.sqrt(2.0 / np.sum(amp ** 2))
    return ks, np.array(ph), amp


def _noise(lin):
    """1e-6 * (hash(global linear index) - 0.5), 24-bit integer hash"""
    h = (lin * 2654435761 + 12345) & 0xFFFFFFFF
    h = ((h ^ (h >> 15)) * 2246822519) & 0xFFFFFFFF
    h = (h ^ (h >> 13)) & 0xFFFFFF
    return 1e-6 * (h.double() / 16777216.0 - 0.5)


def gen_block(kind, gshape, tdtype, device, dim=0, lo=0, hi=None, corner=None):
    """Values of the global field ``kind`` (shape ``gshape``) on the box that is the
    whole grid except [lo, hi) along ``dim``; ``corner`` = c restricts every axis to
    [0, c) instead (a corner of the field, same values).  Generated in fp64, chunk by
    chunk along the first axis of the box."""
    import torch
    D = len(gshape)
    box = [(0, n) for n in gshape]
    if corner is not None:
        box = [(0, min(corner, n)) for n in gshape]
    else:
        box[dim] = (lo, gshape[dim] if hi is None else hi)
    bshape = tuple(b - a for a, b in box)
    out = torch.empty(bshape, dtype=tdtype, device=device)
    strides = [int(np.prod(gshape[d + 1:])) for d in range(D)]
    rest = int(np.prod(bshape[1:])) if D > 1 else 1
    step = max(1, (16 << 20) // max(rest, 1))  # ~128 MB of fp64 per temporary

    def axis(d, a, b, dtype=torch.float64):
        view = [1] * D
        view[d] = b - a
        return torch.arange(a, b, device=device, dtype=dtype).view(view)

    if kind == "nyx":
        ks, ph, amp = _nyx_waves()
        tk = torch.tensor(ks, device=device)
        tph = torch.tensor(ph, device=device)
        tamp = torch.tensor(amp, device=device)

        def table(d, a, b):  # [waves, b - a] complex: exp(i 2 pi k_d x_d)
            x = torch.arange(a, b, device=device, dtype=torch.float64) / gshape[d]
            return torch.polar(torch.ones_like(x)[None, :].expand(len(ks), -1).contiguous(),
                               2.0 * math.pi * tk[:, d:d + 1] * x[None, :])
        Ey = table(1, *box[1])
        Ez = table(2, *box[2]) * torch.polar(tamp, tph)[:, None]
    for a0 in range(box[0][0], box[0][1], step):
        a1 = min(a0 + step, box[0][1])
        lin = axis(0, a0, a1, torch.int64) * strides[0]
        for d in range(1, D):
            lin = lin + axis(d, *box[d], torch.int64) * strides[d]
        if kind == "smooth":
            v = torch.ones((), dtype=torch.float64, device=device)
            ph_ = torch.zeros((), dtype=torch.float64, device=device)
            for d in range(D):
                x = axis(d, a0, a1) if d == 0 else axis(d, *box[d])
                x = x / gshape[d]
                v = v * torch.sin(2.0 * math.pi * x + 0.5 * d)
                ph_ = ph_ + (d + 1) * x
            for k in range(1, 5):
                v = v + 0.05 * 2.0 ** (-k) * torch.sin(2.0 * math.pi * (2 * k + 1) * ph_)
        elif kind == "nyx":
            Ex = table(0, a0, a1)  # [waves, nx]
            # g[i, j, k] = Im sum_m Ex[m, i] Ey[m, j] Ez[m, k]  (one complex GEMM per chunk)
            left = (Ex[:, :, None] * Ey[:, None, :]).reshape(len(ks), -1)  # [m, i*j]
            g = torch.matmul(left.transpose(0, 1), Ez).imag
            v = torch.exp(2.5 * g).reshape(a1 - a0, bshape[1], bshape[2])
        elif kind == "xgc":
            plane = axis(0, a0, a1) * (2.0 * math.pi / gshape[0])
            spec = 1.0 + 0.05 * axis(1, *box[1]) / gshape[1]
            vperp = -4.0 + 8.0 * axis(2, *box[2]) / (gshape[2] - 1)
            vpar = -4.0 + 8.0 * axis(3, *box[3]) / (gshape[3] - 1)
            node = axis(4, *box[4]) / gshape[4]
            v = (torch.exp(-(vpar * vpar + vperp * vperp) / 2.0) * spec
                 * (1.0 + 0.1 * torch.sin(2.0 * math.pi * 7.0 * node + plane)))
        else:
            raise ValueError(kind)
        out[a0 - box[0][0]:a1 - box[0][0]] = (v + _noise(lin)).to(tdtype)
    return out


# ---------------------------------------------------------------------------
# workloads
# ---------------------------------------------------------------------------
class Workload:
    def __init__(self, name, gshape, dtype, s, tol, kind, slab_dim=None, slab_size=None,
                 text=""):
        self.name, self.gshape, self.dtype = name, tuple(gshape), dtype
        self.s, self.tol, self.kind = s, tol, kind
        self.slab_dim, self.slab_size = slab_dim, slab_size
        self.text = text
        self.es = 4 if dtype == "f32" else 8

    def slabs(self):
        """[(lo, hi)] along slab_dim (MaxDim rule: ceil(n / size) slabs of `size` rows,
        the last one shorter)"""
        if self.slab_dim is None:
            return [(0, self.gshape[0])]
        n = self.gshape[self.slab_dim]
        return [(a, min(a + self.slab_size, n)) for a in range(0, n, self.slab_size)]

    def slab_shape(self, lo, hi):
        if self.slab_dim is None:
            return self.gshape
        sh = list(self.gshape)
        sh[self.slab_dim] = hi - lo
        return tuple(sh)

    def elems(self):
        return int(np.prod(self.gshape))


def make_workload(args, world):
    cfg = args.config
    if cfg == "auto":
        cfg = "cfg3" if world == 1 else "cfg5"
    if cfg == "cfg3":
        e = args.edge
        tol = args.tol or 1e-3
        return Workload("cfg3", (e, e, e), "f32", INF, tol, args.field, None, None,
                        "3D fp32 %d^3 %s synthetic field%s, rel L_inf %g, compress (Huffman, "
                        "dict 8192, chunk 1024)" % (
                            e, "NYX-shaped (exp(2.5 g), 32 plane waves)" if args.field == "nyx"
                            else "smooth", " per GPU" if world > 1 else "", tol))
    if cfg == "cfg5":
        tol = args.tol or 1e-4
        return Workload("cfg5", (2049, 2049, 2049), "f32", INF, tol, "smooth", 0, 257,
                        "3D fp32 2049^3 smooth synthetic field, rel L_inf %g, MaxDim "
                        "subdomains 7 x 257 + 250 rows shared over the GPUs" % tol)
    tol = args.tol or 1e-3
    return Workload("cfg4", (8, 16, 39, 39, 39969), "f64", 1.0, tol, "xgc", 4, 4997,
                    "5D fp64 8x16x39x39x39969 XGC-shaped synthetic field, rel s-norm (s=1) %g, "
                    "8 subdomains along the last axis (7 x 4997 + 4990)" % tol)


# ---------------------------------------------------------------------------
# clocks sampler (nvidia-smi during the timed region)
# ---------------------------------------------------------------------------
class Clocks:
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index = index
        self.rows = []
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                 "--format=csv,noheader,nounits", "-lms", "100"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown",
                 "sw_power_cap"]
        for r in self.rows:
            if len(r) < 6:
                continue
            try:
                sm.append(float(r[0]))
                mx.append(float(r[1]))
            except ValueError:
                continue
            for n, v in zip(names, r[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": float(np.median(sm)) if sm else None,
                "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


# ---------------------------------------------------------------------------
# CPU arm: the reference's OPENMP Compressor (oracle/_ref) or the C port
# ---------------------------------------------------------------------------
REF_SO = os.path.join(ROOT, "oracle", "_ref", "libmgardx_ref.so")
PORT_SO = os.path.join(ROOT, "oracle", "libmgx_oracle.so")


def host_cores():
    n = len(os.sched_getaffinity(0))
    env = os.environ.get("OMP_NUM_THREADS")
    if env and env.isdigit():
        n = min(n, int(env))
    return n


def cpu_run(arr, tol, s, reps):
    """reps x Compress and reps x Decompress of ``arr`` (numpy, 1-5 D, fp32/fp64) with the
    relative bound ``tol`` in the s-norm.  Returns per-rep seconds and what ran."""
    dt = arr.dtype
    arr = np.ascontiguousarray(arr)
    D = arr.ndim
    shape = (ctypes.c_uint64 * D)(*arr.shape)
    cs = (ctypes.c_double * reps)()
    ds = (ctypes.c_double * reps)()
    nb = ctypes.c_uint64()
    err = ctypes.c_double()
    tname = "float" if dt == np.float32 else "double"
    if os.path.exists(REF_SO):
        lib = ctypes.CDLL(REF_SO)
        fn = getattr(lib, "ref_bench_%d%s" % (D, "f" if dt == np.float32 else "d"))
        rc = fn(1, shape, arr.ctypes.data_as(ctypes.c_void_p), 1,
                ctypes.c_double(tol), ctypes.c_double(s), reps, cs, ds,
                ctypes.byref(nb), ctypes.byref(err))
        if rc != 0:
            raise RuntimeError("reference bench failed rc=%d" % rc)
        kind, cores = "reference", host_cores()
        what = "reference mgard_x::Compressor<%d,%s,OPENMP> (oracle/_ref)" % (D, tname)
    else:
        if not os.path.exists(PORT_SO):
            subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle"), "port"],
                                  stdout=subprocess.DEVNULL)
        lib = ctypes.CDLL(PORT_SO)
        suf = "f32" if dt == np.float32 else "f64"
        cap = arr.nbytes * 3 + (1 << 20)
        out = np.zeros(cap, np.uint8)
        back = np.zeros_like(arr)
        nrm = ctypes.c_double(0.0)
        for r in range(reps):
            nrm = ctypes.c_double(0.0)
            t0 = time.perf_counter()
            rc = getattr(lib, "mgxo_compress_" + suf)(
                D, shape, arr.ctypes.data_as(ctypes.c_void_p), 1, ctypes.c_double(tol),
                ctypes.c_double(s), ctypes.byref(nrm), 8192, 1024,
                out.ctypes.data_as(ctypes.c_void_p), ctypes.c_uint64(cap), ctypes.byref(nb))
            cs[r] = time.perf_counter() - t0
            assert rc == 0
        for r in range(reps):
            t0 = time.perf_counter()
            rc = getattr(lib, "mgxo_decompress_" + suf)(
                D, shape, out.ctypes.data_as(ctypes.c_void_p), ctypes.c_uint64(nb.value), 1,
                ctypes.c_double(tol), ctypes.c_double(s), nrm, 8192,
                back.ctypes.data_as(ctypes.c_void_p))
            ds[r] = time.perf_counter() - t0
            assert rc == 0
        err.value = float(np.abs(back.astype(np.float64) - arr).max())
        kind, cores = "port", 1
        what = "C restatement oracle/mgx_oracle.c (scalar)"
    return {"compress_s": list(cs), "decompress_s": list(ds), "bytes": nb.value,
            "max_abs_err": err.value, "kind": kind, "cores": cores, "what": what,
            "in_bytes": arr.nbytes}


def gen_host(kind, gshape, dtype, corner=None, dim=0, lo=0, hi=None):
    """numpy copy of a field box for the CPU arm (generated on the GPU when there is one:
    the generator is torch code, not a kernel of this repository)"""
    import torch
    dev = "cuda" if torch.cuda.is_available() else "cpu"
    tdt = torch.float32 if dtype == "f32" else torch.float64
    return gen_block(kind, gshape, tdt, dev, dim, lo, hi, corner).cpu().numpy()


def cpu_sample(wl, budget_s, reps=1, what=None):
    """CPU arm on the largest sample of ``wl`` that fits ``budget_s`` seconds of work per
    rep: the whole field (or one whole subdomain), else a corner / shorter slab of it.
    Returns (cpu_run result, description of the sample)."""
    # calibrate on a small corner: seconds per input byte (compress + decompress)
    D = len(wl.gshape)
    if wl.slab_dim is None:
        full_shape, cal_corner = wl.gshape, (129 if D == 3 else 1025 if D == 2 else 4097)
        cal = cpu_run(gen_host(wl.kind, wl.gshape, wl.dtype, corner=cal_corner), wl.tol, wl.s, 1)
        per_byte = (cal["compress_s"][-1] + cal["decompress_s"][-1]) / cal["in_bytes"]
        full_bytes = wl.elems() * wl.es
        if per_byte * full_bytes <= budget_s:
            arr = gen_host(wl.kind, wl.gshape, wl.dtype)
            desc = "the whole %s field (%.0f MB)" % ("x".join(map(str, wl.gshape)),
                                                     full_bytes / 1e6)
            r = cpu_run(arr, wl.tol, wl.s, reps)
            return r, desc, True
        c = cal_corner
        while True:
            nxt = 2 * (c - 1) + 1
            if nxt > min(wl.gshape) or per_byte * (nxt ** D) * wl.es > budget_s:
                break
            c = nxt
        arr = gen_host(wl.kind, wl.gshape, wl.dtype, corner=c)
        desc = "the [0,%d)^%d corner of the %s field (%.0f MB)" % (
            c, D, "x".join(map(str, wl.gshape)), arr.nbytes / 1e6)
        r = cal if c == cal_corner and reps == 1 else cpu_run(arr, wl.tol, wl.s, reps)
        return r, desc, False
    # decomposed workload: a (possibly shortened) first subdomain
    lo, hi = wl.slabs()[0]
    n = hi - lo
    short = max(9, min(n, 33))
    cal = cpu_run(gen_host(wl.kind, wl.gshape, wl.dtype, dim=wl.slab_dim, lo=0, hi=short),
                  wl.tol, wl.s, 1)
    per_byte = (cal["compress_s"][-1] + cal["decompress_s"][-1]) / cal["in_bytes"]
    rows = short
    while rows < n and per_byte * (cal["in_bytes"] / short) * min(n, 2 * (rows - 1) + 1) <= budget_s:
        rows = min(n, 2 * (rows - 1) + 1)
    if rows == short and reps == 1:
        r = cal
    else:
        r = cpu_run(gen_host(wl.kind, wl.gshape, wl.dtype, dim=wl.slab_dim, lo=0, hi=rows),
                    wl.tol, wl.s, reps)
    sh = list(wl.gshape)
    sh[wl.slab_dim] = rows
    desc = "%s of subdomain 0 (%s, %.0f MB; a subdomain has %d rows along axis %d)" % (
        "the whole" if rows == n else "the first %d rows" % rows,
        "x".join(map(str, sh)), r["in_bytes"] / 1e6, n, wl.slab_dim)
    return r, desc, rows == n


def hbm_peak():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


def dram_traffic(stage):
    """per-launch DRAM bytes of the stage's kernel from the committed
    ncu --set full capture (profiles/dram_traffic.json), or None."""
    try:
        with open(os.path.join(ROOT, "profiles", "dram_traffic.json")) as f:
            return json.load(f).get(stage)
    except Exception:
        return None


def cpu_baseline_leg(wl, budget_s):
    r, desc, whole = cpu_sample(wl, budget_s)
    gbs = r["in_bytes"] / min(r["compress_s"]) / 1e9
    return {"value": gbs, "unit": UNIT, "cores": r["cores"], "kind": r["kind"],
            "sample": "%s, one Compress of %s, rel bound %g (s = %s); decompress %.3f GB/s; "
                      "CR %.2f" % (r["what"], desc, wl.tol, wl.s,
                                   r["in_bytes"] / min(r["decompress_s"]) / 1e9,
                                   r["in_bytes"] / max(r["bytes"], 1)),
            "whole_workload": whole,
            "decompress": r["in_bytes"] / min(r["decompress_s"]) / 1e9}


# ---------------------------------------------------------------------------
# stage model: algorithmic bytes of this implementation's passes
# ---------------------------------------------------------------------------
C_STAGES = [("norm", "c_begin", "c_norm"), ("refactor_quantize", "c_norm", "c_refactor"),
            ("codebook", "c_refactor", "c_codebook"), ("encode", "c_codebook", "c_sizes"),
            ("pack", "c_sizes", "c_pack"), ("outliers_tail", "c_pack", "c_end")]
D_STAGES = [("decode", "d_begin", "d_decode"), ("dequantize", "d_decode", "d_dequant"),
            ("recompose", "d_dequant", "d_end")]
STAGE_KERNELS = {"norm": "k_norm", "refactor_quantize": "k_fwd3 (+ k_fwd3_rest, k_hist_fix) "
                 "or the level kernels of the orthogonal path", "codebook": "k_codebook",
                 "encode": "k_encode", "pack": "k_pack", "outliers_tail": "outlier sort + header"}


def roofline_of(stage_ms, n_elems, es, comp_bytes, step_ms, peak, peak_src, rel=True,
                traffic_ok=False):
    algo = {"norm": float(es * n_elems) if rel else 0.0,
            "refactor_quantize": float((es + 2) * n_elems), "codebook": 0.0,
            "encode": 2.0 * n_elems + comp_bytes, "pack": 0.0, "outliers_tail": 0.0}
    dom = max(stage_ms, key=lambda k: stage_ms[k])
    dom_ms = stage_ms[dom]
    achieved = algo[dom] / (dom_ms * 1e-3) / 1e9 if dom_ms > 0 else 0.0
    strict = (es * n_elems + comp_bytes) / (step_ms * 1e-3) / 1e9
    passes = sum(algo.values()) / (step_ms * 1e-3) / 1e9
    return {"bound": "hbm", "kernel": dom, "kernels_of_stage": STAGE_KERNELS.get(dom),
            "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
            "traffic": dram_traffic(dom) if traffic_ok else None, "peak_source": peak_src,
            "algorithmic_bytes_per_launch": algo[dom], "kernel_ms": dom_ms,
            "step_frac": strict / peak,
            "step_frac_definition": "(N*sizeof(T) + compressed bytes) / step time / peak "
                                    "(SURVEY 8d)",
            "pass_model_frac": passes / peak,
            "stages_ms": dict(stage_ms)}


# ---------------------------------------------------------------------------
# the other BASELINE configurations (N = 1; reported, not the metric)
# ---------------------------------------------------------------------------
def short_run(wl, dev, peak, peak_src, reps=3, lo=None, hi=None, eb_abs=None):
    """median-of-reps compress / decompress of one field (or one subdomain of a decomposed
    workload) with stage times, strict roofline fraction and the error check"""
    import torch
    import mgard_b200 as mg
    tdt = torch.float32 if wl.dtype == "f32" else torch.float64
    if wl.slab_dim is None:
        u = gen_block(wl.kind, wl.gshape, tdt, dev)
    else:
        u = gen_block(wl.kind, wl.gshape, tdt, dev, wl.slab_dim, lo, hi)
    shape = tuple(u.shape)
    comp = mg.Compressor(mg.Hierarchy(shape, tdt, mg.Config(dev_id=dev.index or 0)))
    comp.set_timing(True)
    buf = torch.empty(u.numel() * wl.es + (1 << 20), dtype=torch.uint8, device=dev)
    back = torch.empty_like(u)
    eb, tol = (mg.REL, wl.tol) if eb_abs is None else (mg.ABS, eb_abs)
    cs, ds, st, dst = [], [], [], []
    l0 = comp.launch_count()
    for it in range(reps + 1):
        blob, norm = comp.Compress(u, eb, tol, wl.s, out=buf)
        c = comp.elapsed_ms("c_begin", "c_end")
        s1 = {k: comp.elapsed_ms(a, b) for k, a, b in C_STAGES}
        comp.Decompress(blob, eb, tol, wl.s, norm, out=back)
        d = comp.elapsed_ms("d_begin", "d_end")
        s2 = {k: comp.elapsed_ms(a, b) for k, a, b in D_STAGES}
        if it:
            cs.append(c)
            ds.append(d)
            st.append(s1)
            dst.append(s2)
    launches = (comp.launch_count() - l0) // (reps + 1)
    nbytes = u.numel() * wl.es
    cms, dms = float(np.median(cs)), float(np.median(ds))
    stage_ms = {k: float(np.median([x[k] for x in st])) for k, _, _ in C_STAGES}
    dstage_ms = {k: float(np.median([x[k] for x in dst])) for k, _, _ in D_STAGES}
    diff = back.double() - u.double()
    if math.isinf(wl.s):
        bound = tol * (norm if eb == mg.REL else 1.0)
        err = float(diff.abs().max().item()) / bound
    else:
        l2 = math.sqrt(float((diff * diff).sum().item()) / u.numel())
        err = l2 / (tol * (norm if eb == mg.REL else 1.0))
    rl = roofline_of(stage_ms, u.numel(), wl.es, int(blob.numel()), cms, peak, peak_src,
                     rel=eb == mg.REL)
    rl["decompress_stages_ms"] = dstage_ms
    out = {"workload": wl.text if wl.slab_dim is None else
           "%s -- subdomain %s" % (wl.text, "x".join(map(str, shape))),
           "compress_ms": cms, "decompress_ms": dms,
           "compress_GBs": nbytes / cms / 1e6, "decompress_GBs": nbytes / dms / 1e6,
           "compression_ratio": nbytes / int(blob.numel()),
           "err_over_bound": err, "launches_per_pair": int(launches), "roofline": rl}
    del comp, u, buf, back
    torch.cuda.empty_cache()
    return out


def other_configs(args, dev, peak, peak_src, with_cpu):
    """cfg 3 at the other end of the tolerance sweep and on the smooth field, cfg 1, cfg 2,
    one subdomain of cfg 4 and of cfg 5"""
    e = args.edge
    items = [
        ("cfg3_nyx_1e-5", Workload("cfg3", (e, e, e), "f32", INF, 1e-5, "nyx", text=(
            "3D fp32 %d^3 NYX-shaped synthetic field, rel L_inf 1e-05" % e)), {}, 0.0),
        ("cfg3_smooth_1e-3", Workload("cfg3s", (e, e, e), "f32", INF, 1e-3, "smooth", text=(
            "3D fp32 %d^3 smooth synthetic field, rel L_inf 0.001 (round 1's field family)"
            % e)), {}, 0.0),
        ("cfg1_513^3", Workload("cfg1", (513, 513, 513), "f32", INF, 1e-3, "smooth", text=(
            "3D fp32 513^3 smooth synthetic field, rel L_inf 0.001")), {}, 8.0),
        ("cfg2_8193^2", Workload("cfg2", (8193, 8193), "f64", 0.0, 1e-4, "smooth", text=(
            "2D fp64 8193^2 smooth synthetic field, rel L2 (s=0) 0.0001")), {}, 12.0),
        ("cfg4_subdomain", Workload("cfg4", (8, 16, 39, 39, 39969), "f64", 1.0, 1e-3, "xgc", 4,
                                    4997, "5D fp64 8x16x39x39x39969 XGC-shaped field, rel "
                                          "s-norm (s=1) 0.001"), {"lo": 0, "hi": 4997}, 10.0),
        ("cfg5_subdomain", Workload("cfg5", (2049, 2049, 2049), "f32", INF, 1e-4, "smooth", 0,
                                    257, "3D fp32 2049^3 smooth field, rel L_inf 0.0001"),
         {"lo": 0, "hi": 257}, 6.0),
    ]
    out = {}
    for name, wl, kw, cpu_budget in items:
        try:
            r = short_run(wl, dev, peak, peak_src, **kw)
            if with_cpu and cpu_budget > 0:
                r["cpu_baseline"] = cpu_baseline_leg(wl, cpu_budget)
            out[name] = r
        except Exception as ex:  # reported, never fatal for the metric line
            out[name] = {"error": repr(ex)}
    return out


# ---------------------------------------------------------------------------
# --impl reference: the reference's CPU implementation of the path
# ---------------------------------------------------------------------------
def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if rank != 0:
        return
    # torchrun pins OMP_NUM_THREADS=1 for its workers; the CPU arm is meant to
    # use every host core it can (set before the OpenMP runtime is loaded)
    os.environ["OMP_NUM_THREADS"] = str(len(os.sched_getaffinity(0)))
    wl = make_workload(args, max(world, args.gpus))
    reps = args.steps + args.warmup
    budget = 150.0 / max(reps, 1)
    r, desc, whole = cpu_sample(wl, budget, reps=reps)
    t = sum(r["compress_s"][args.warmup:])
    gbs = args.steps * r["in_bytes"] / t / 1e9
    workload = wl.text if whole and wl.slab_dim is None else wl.text + " -- CPU SAMPLE: " + desc
    line = {
        "impl": "reference", "metric": METRIC, "value": gbs, "unit": UNIT,
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": 1e3 * t / args.steps, "higher_is_better": True,
        "scaling": "weak" if wl.slab_dim is None else "strong", "vs_baseline": None,
        "dtype": wl.dtype, "data": "synthetic",
        "config": {"workload": workload},
        "cpu_baseline": {"value": gbs, "unit": UNIT, "cores": r["cores"], "kind": r["kind"],
                         "sample": "%s; each step = one Compress of %s" % (r["what"], desc),
                         "whole_workload": bool(whole and wl.slab_dim is None)},
        "e2e": {"value": gbs, "unit": UNIT, "h2d_bytes_per_step": 0,
                "d2h_bytes_per_step": 0},
        "decompress": {"value": args.steps * r["in_bytes"]
                       / sum(r["decompress_s"][args.warmup:]) / 1e9, "unit": UNIT},
        "compression_ratio": r["in_bytes"] / max(r["bytes"], 1),
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ---------------------------------------------------------------------------
# N ranks -> one "MGARD" stream -> mgx_x_decompress on one GPU
# ---------------------------------------------------------------------------
def check_stream(dom, blobs, owned, slabs, wl, tdt, dev, local, rank, norm, ebtype):
    """Rank 0 gathers every rank's records behind one MGARD header (what
    mgard_x::compress returns for the whole domain under the MaxDim decomposition),
    decodes that stream ALONE through mgx_x_decompress and checks the bound on every
    element against the regenerated field."""
    import psutil
    import torch
    from mgard_b200 import highlevel as hl
    t0 = time.perf_counter()
    data = dom.gather_stream(blobs, owned, wl.gshape, ebtype, wl.tol, wl.s, norm,
                             wl.slab_dim, wl.slab_size)
    if rank != 0:
        return None
    field_bytes = wl.elems() * wl.es
    out = {"stream_bytes": len(data), "records": len(slabs),
           "compression_ratio": field_bytes / len(data),
           "gather_s": time.perf_counter() - t0}
    avail = psutil.virtual_memory().available
    if avail < 2 * field_bytes + (16 << 30):
        out["decoded"] = False
        out["why"] = "host has %.0f GB free, the decoded field needs %.0f GB" % (
            avail / 1e9, field_bytes / 1e9)
        return out
    t0 = time.perf_counter()
    back = hl.decompress(np.frombuffer(data, np.uint8), config=hl.Config(dev_id=local))
    out["decode_s"] = time.perf_counter() - t0
    assert tuple(back.shape) == tuple(wl.gshape)
    worst, sq = 0.0, 0.0
    for lo, hi in slabs:
        sl = [slice(None)] * len(wl.gshape)
        sl[wl.slab_dim] = slice(lo, hi)
        ref = gen_block(wl.kind, wl.gshape, tdt, dev, wl.slab_dim, lo, hi)
        got = torch.from_numpy(np.ascontiguousarray(back[tuple(sl)])).to(dev)
        d = got.double() - ref.double()
        worst = max(worst, float(d.abs().max().item()))
        sq += float((d * d).sum().item())
        del ref, got, d
    if math.isinf(wl.s):
        ratio = worst / (wl.tol * norm)
    else:
        ratio = math.sqrt(sq / wl.elems()) / (wl.tol * norm)
    out.update({"decoded": True, "decoder": "mgx_x_decompress on rank 0's GPU alone",
                "err_over_bound": ratio, "ok": bool(ratio <= 1.0),
                "checked_elements": wl.elems()})
    return out


# ---------------------------------------------------------------------------
# our arm
# ---------------------------------------------------------------------------
def run_ours(args):
    import torch
    import torch.distributed as dist
    import mgard_b200 as mg
    from mgard_b200.sharding import ShardedDomain

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world == 1 and args.gpus > 1:
        # plain "python bench.py --gpus N": re-launch one rank per GPU
        port = 29500 + os.getpid() % 2000
        os.execv(sys.executable, [
            sys.executable, "-m", "torch.distributed.run", "--nnodes=1",
            "--nproc-per-node", str(args.gpus), "--master-addr", "127.0.0.1",
            "--master-port", str(port)] + sys.argv)
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (there is no CPU fallback)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    numa = None
    if world > 1 and not args.no_numa:
        from mgard_b200.sharding import bind_to_gpu_numa
        numa = bind_to_gpu_numa(local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        # rank 0's stdout carries ONE JSON line: keep NCCL's version banner off it
        if os.environ.get("NCCL_DEBUG", "").upper() in ("", "VERSION"):
            os.environ["NCCL_DEBUG"] = "WARN"
        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    wl = make_workload(args, world)
    K, W = args.steps, max(args.warmup, 0)
    tdt = torch.float32 if wl.dtype == "f32" else torch.float64
    stream = torch.cuda.Stream(device=dev)
    weak = wl.slab_dim is None  # one whole field per rank
    if weak:
        # cfg3: every rank its own cube (ranks > 0 see a scaled copy so that the global
        # value range differs from the local ones)
        slabs = [(0, wl.gshape[0])]
        owned = [0]
        fields = [gen_block(wl.kind, wl.gshape, tdt, dev) * (1.0 + 0.25 * rank / world)]
        total_elems = wl.elems() * world
        n_sub = world
    else:
        slabs = wl.slabs()
        per = -(-len(slabs) // world)
        owned = [i for i in range(len(slabs)) if i // per == rank]
        fields = [gen_block(wl.kind, wl.gshape, tdt, dev, wl.slab_dim, *slabs[i]) for i in owned]
        total_elems = wl.elems()
        n_sub = len(slabs)
    dom = ShardedDomain([tuple(f.shape) for f in fields], tdt, total_elems, n_sub,
                        device=dev, stream=stream, dev_id=local)
    my_elems = sum(f.numel() for f in fields)
    my_bytes = my_elems * wl.es
    all_bytes = torch.tensor([float(my_bytes)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(all_bytes)
    job_bytes = float(all_bytes.item())
    outs = [torch.empty(f.numel() * wl.es // 2 + (8 << 20), dtype=torch.uint8, device=dev)
            for f in fields]
    state = {}

    def compress_step():
        blobs, norm, ltol = dom.Compress(fields, mg.REL, wl.tol, wl.s, outs=outs)
        state["blobs"], state["norm"], state["ltol"] = blobs, norm, ltol

    def timed(fn, steps):
        barrier()
        e0 = torch.cuda.Event(enable_timing=True)
        e1 = torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        e0.record(stream)
        for _ in range(steps):
            fn()
        e1.record(stream)
        barrier()
        wall = (time.perf_counter() - t0) * 1e3
        ms = torch.tensor([e0.elapsed_time(e1), wall], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms[0].item()), float(ms[1].item())

    # ---- compress, inputs resident in HBM ------------------------------------
    for _ in range(W):
        compress_step()
    dom.set_timing(True)
    stage_ms = {k: 0.0 for k, _, _ in C_STAGES}

    def compress_step_prof():
        compress_step()
        for k, v in dom.stage_ms(C_STAGES).items():
            stage_ms[k] += v
        stage_ms["norm"] += dom.last_norm_ms

    clocks = Clocks(local)
    if rank == 0:
        clocks.start()
    l0 = dom.launch_count()
    ms, wall = timed(compress_step_prof, K)
    launches = dom.launch_count() - l0
    clk = clocks.stop() if rank == 0 else None
    dom.set_timing(False)
    blobs, norm, ltol = state["blobs"], state["norm"], state["ltol"]
    comp_bytes = sum(int(b.numel()) for b in blobs)
    value = job_bytes * K / (ms * 1e-3) / 1e9

    # ---- decompress, resident -------------------------------------------------
    keep = [b.clone() for b in blobs]
    backs = [torch.empty_like(f) for f in fields]

    def decompress_step():
        dom.Decompress(keep, ltol, wl.s, outs=backs)

    for _ in range(W):
        decompress_step()
    dom.set_timing(True)
    dstage_ms = {k: 0.0 for k, _, _ in D_STAGES}

    def decompress_step_prof():
        decompress_step()
        for k, v in dom.stage_ms(D_STAGES).items():
            dstage_ms[k] += v

    dms, _ = timed(decompress_step_prof, K)
    dom.set_timing(False)
    dvalue = job_bytes * K / (dms * 1e-3) / 1e9
    # error bound on EVERY element (outside the timed regions)
    if math.isinf(wl.s):
        worst = max(float((b.double() - f.double()).abs().max().item())
                    for b, f in zip(backs, fields))
        eb = torch.tensor([worst / (wl.tol * norm)], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(eb, op=dist.ReduceOp.MAX)
        eb_ratio = float(eb.item())
    else:
        sq = sum(float(((b.double() - f.double()) ** 2).sum().item())
                 for b, f in zip(backs, fields))
        t = torch.tensor([sq], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t)
        eb_ratio = math.sqrt(float(t.item()) / total_elems) / (wl.tol * norm)
    del backs

    # ---- end to end through the host-pointer call ------------------------------
    e2e = None
    if not args.no_e2e:
        h_in = [torch.empty(f.shape, dtype=tdt, pin_memory=True) for f in fields]
        for h, f in zip(h_in, fields):
            h.copy_(f)
        h_out = [torch.empty(f.numel() * wl.es // 2 + (8 << 20), dtype=torch.uint8,
                             pin_memory=True) for f in fields]
        h_back = [torch.empty(f.shape, dtype=tdt, pin_memory=True) for f in fields]
        a_in = [h.numpy() for h in h_in]
        a_out = [h.numpy() for h in h_out]
        a_back = [h.numpy() for h in h_back]
        from mgard_b200 import highlevel as hl
        hcfg = hl.Config(dev_id=local)
        single = world == 1 and weak

        def e2e_step():
            if single:
                # the call a user of the reference makes: mgard_x::compress with host
                # pointers -> self-describing "MGARD" stream
                state["hblobs"] = [hl.compress(a_in[0], wl.tol, wl.s, hl.REL, config=hcfg,
                                               out=a_out[0])]
            else:
                # sharded: global norm first (device partials of the slabs just copied),
                # then one host-pointer compress per subdomain with the ABS bound
                state["hblobs"] = dom.CompressHost(a_in, mg.REL, wl.tol, wl.s, outs=a_out)

        def e2e_dstep():
            if single:
                hl.decompress(state["hblobs"][0], config=hcfg, out=a_back[0])
            else:
                dom.DecompressHost(state["hblobs"], ltol, wl.s, outs=a_back)

        for _ in range(max(min(W, 2), 1)):
            e2e_step()
        ems, ewall = timed(e2e_step, K)
        for _ in range(max(min(W, 2), 1)):
            e2e_dstep()
        edms, edwall = timed(e2e_dstep, K)
        hbytes = torch.tensor([float(sum(int(b.size) for b in state["hblobs"]))],
                              dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(hbytes)
        e2e = {"value": job_bytes * K / (max(ems, ewall) * 1e-3) / 1e9,
               "unit": UNIT, "h2d_bytes_per_step": int(job_bytes),
               "d2h_bytes_per_step": int(hbytes.item()),
               "h2d_GBs_per_gpu": my_bytes * K / (max(ems, ewall) * 1e-3) / 1e9,
               "decompress": job_bytes * K / (max(edms, edwall) * 1e-3) / 1e9,
               "rank0_numa_binding": numa,
               "note": "pinned host buffers; wall clock (>= device time) over K calls of "
                       + ("mgx_x_compress / mgx_x_decompress (= mgard_x::compress / "
                          "decompress: MGARD header + records)" if single else
                          "mgx_norm_partial + all-reduce + mgx_compress_host per subdomain "
                          "/ mgx_decompress_host") + ", max over ranks"}
        if single:
            e2e["max_err_over_bound"] = float(
                np.abs(a_back[0].astype(np.float64) - a_in[0].astype(np.float64)).max()
                / (wl.tol * norm))
        del h_in, h_out, h_back

    # ---- N ranks -> ONE stream, decoded by ONE GPU (outside the timed regions) ----------
    stream_check = None
    if world > 1 and not weak and not args.no_stream_check:
        stream_check = check_stream(dom, blobs, owned, slabs, wl, tdt, dev, local, rank,
                                    norm, mg.REL)

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    # ---- roofline ---------------------------------------------------------------
    peak, peak_src = hbm_peak()
    stage_avg = {k: v / K for k, v in stage_ms.items()}
    roofline = roofline_of(stage_avg, my_elems, wl.es, comp_bytes, ms / K, peak, peak_src,
                           traffic_ok=(wl.name == "cfg3" and wl.gshape == (1025, 1025, 1025)))
    roofline["decompress_stages_ms"] = {k: v / K for k, v in dstage_ms.items()}
    roofline["note"] = "rank 0's subdomains; stage times are CUDA events inside the library"
    cpu = None
    if not args.no_cpu and world == 1:
        cpu = cpu_baseline_leg(wl, 25.0)
    extra = None
    if not args.no_extra and world == 1 and weak:
        del fields, outs, keep
        torch.cuda.empty_cache()
        extra = other_configs(args, dev, peak, peak_src, not args.no_cpu)
    sharding = "single GPU"
    if world > 1 and weak:
        sharding = "one %s field per GPU, NCCL allreduce(MAX) of the value range only" % (
            "x".join(map(str, wl.gshape)))
    elif not weak:
        rows = [hi - lo for lo, hi in slabs]
        sharding = ("%d MaxDim subdomains along axis %d (%s rows), contiguous blocks of %d per "
                    "rank (rank 0: %s); one NCCL allreduce(%s) of one fp64 scalar; "
                    "largest / mean rank load %.3f" % (
                        len(slabs), wl.slab_dim, "+".join(map(str, sorted(set(rows), reverse=True))),
                        -(-len(slabs) // world), owned, "MAX" if math.isinf(wl.s) else "SUM",
                        max(sum(rows[i] for i in range(len(slabs)) if i // (-(-len(slabs) // world)) == r)
                            for r in range(world)) / (sum(rows) / world)))
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world,
        "steps": K, "warmup": W, "ms_per_step": ms / K, "higher_is_better": True,
        "scaling": "weak" if weak else "strong", "vs_baseline": None, "dtype": wl.dtype,
        "data": "synthetic",
        "config": {"workload": wl.text,
                   "l2": "input %.2f GB per rank and step >> %d MB L2 (no flush needed)" % (
                       my_bytes / 1e9, L2_BYTES >> 20),
                   "sharding": sharding},
        "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e, "clocks": clk,
        "gpu_launches": int(launches),
        "decompress": {"value": dvalue, "unit": UNIT, "ms_per_step": dms / K},
        "compression_ratio": job_bytes / max(comp_bytes * (world if weak else 1), 1)
        if weak else None,
        "rank0_compression_ratio": my_bytes / max(comp_bytes, 1),
        "error_bound": {"max_err_over_bound": eb_ratio, "ok": bool(eb_ratio <= 1.0),
                        "checked_elements": int(total_elems)},
        "wall_ms_per_step": wall / K,
        "stream_check": stream_check,
        "other_configs": extra,
    }
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    a = parse()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_ours(a)
