"""Stage timing of one configuration through the low-level Compressor mirror
(device-resident input), e.g. BASELINE config 2:

  python tools/stage_times.py --shape 8193 8193 --dtype f64 --s 0 --tol 1e-4

Not part of the bench contract (bench.py measures config 3); this is the
measuring stick for the correction (orthogonal projection) path."""
import argparse
import json
import math
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import mgard_b200 as mg  # noqa: E402


def field(shape, dtype, device):
    """smooth separable product + a skewed plane wave + tiny hash noise"""
    D = len(shape)
    v = torch.ones(shape, dtype=torch.float64, device=device)
    ph = torch.zeros(shape, dtype=torch.float64, device=device)
    for d, n in enumerate(shape):
        view = [1] * D
        view[d] = n
        x = (torch.arange(n, device=device, dtype=torch.float64) / n).view(view)
        v = v * torch.sin(2 * math.pi * x + 0.5 * d)
        ph = ph + (d + 1) * x
    for k in range(1, 5):
        v = v + 0.05 * 2.0 ** (-k) * torch.sin(2 * math.pi * (2 * k + 1) * ph)
    lin = torch.arange(v.numel(), device=device, dtype=torch.int64).view(shape)
    h = (lin * 2654435761 + 12345) & 0xFFFFFFFF
    h = (h ^ (h >> 15)) * 2246822519 & 0xFFFFFFFF
    h = (h ^ (h >> 13)) & 0xFFFFFF
    v = v + 1e-6 * (h.to(torch.float64) / 16777216.0 - 0.5)
    return v.to(dtype)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--shape", type=int, nargs="+", required=True)
    ap.add_argument("--dtype", default="f32", choices=["f32", "f64"])
    ap.add_argument("--s", type=float, default=float("inf"))
    ap.add_argument("--tol", type=float, default=1e-3)
    ap.add_argument("--abs", action="store_true")
    ap.add_argument("--field", default="smooth", choices=["smooth", "nyx"],
                    help="nyx: bench.py's NYX-shaped generator (3-D only)")
    ap.add_argument("--compress-only", action="store_true",
                    help="diagnostic builds whose streams do not decode")
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=2)
    a = ap.parse_args()
    dt = torch.float32 if a.dtype == "f32" else torch.float64
    shape = tuple(a.shape)
    dev = torch.device("cuda", 0)
    if a.field == "nyx":
        import bench
        u = bench.gen_block("nyx", shape, dt, dev)
    else:
        u = field(shape, dt, dev)
    comp = mg.Compressor(mg.Hierarchy(shape, dt))
    comp.set_timing(True)
    eb = mg.ABS if a.abs else mg.REL
    out = torch.empty(u.numel() * u.element_size() + (1 << 20), dtype=torch.uint8,
                      device=dev)
    back = torch.empty_like(u)
    cs, ds = [], []
    stages = {}
    for it in range(a.warmup + a.steps):
        blob, norm = comp.Compress(u, eb, a.tol, a.s, out=out)
        c = comp.elapsed_ms("c_begin", "c_end")
        st = {"norm": comp.elapsed_ms("c_begin", "c_norm"),
              "refactor_quantize": comp.elapsed_ms("c_norm", "c_refactor"),
              "codebook": comp.elapsed_ms("c_refactor", "c_codebook"),
              "encode": comp.elapsed_ms("c_codebook", "c_end")}
        d = 0.0
        if not a.compress_only:
            comp.Decompress(blob, eb, a.tol, a.s, norm, out=back)
            d = comp.elapsed_ms("d_begin", "d_end")
            st["decode"] = comp.elapsed_ms("d_begin", "d_decode")
            st["dequant_recompose"] = comp.elapsed_ms("d_decode", "d_end")
        if it >= a.warmup:
            cs.append(c)
            ds.append(d)
            for k, v in st.items():
                stages.setdefault(k, []).append(v)
    nbytes = u.numel() * u.element_size()
    diff = (back.double() - u.double())
    if math.isinf(a.s):
        err = diff.abs().max().item() / (1.0 if a.abs else u.abs().max().item())
    else:
        err = math.sqrt((diff * diff).sum().item() / u.numel())
        if not a.abs:
            err /= math.sqrt((u.double() ** 2).sum().item() / u.numel())
    print(json.dumps({
        "shape": shape, "dtype": a.dtype, "s": a.s, "tol": a.tol,
        "mode": "ABS" if a.abs else "REL",
        "compress_ms": float(np.median(cs)), "decompress_ms": float(np.median(ds)),
        "compress_GBs": nbytes / np.median(cs) / 1e6,
        "decompress_GBs": nbytes / np.median(ds) / 1e6,
        "CR": nbytes / blob.numel(), "err_over_tol": err / a.tol,
        "stages_ms": {k: float(np.median(v)) for k, v in stages.items()},
        "launches": comp.launch_count()}))


if __name__ == "__main__":
    main()
