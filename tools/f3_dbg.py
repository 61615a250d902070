"""k_fwd3 with its compute or its loads switched off (MGX_F3_DBG, read at every launch):
how long the tile pipeline takes when only one side works.  Streams are not valid."""
import os, sys, json, numpy as np, torch
sys.path.insert(0, ".")
import bench, mgard_b200 as mg
shape = (1025, 1025, 1025)
tol = float(sys.argv[1]) if len(sys.argv) > 1 else 1e-3
u = bench.gen_block("nyx", shape, torch.float32, torch.device("cuda", 0))
comp = mg.Compressor(mg.Hierarchy(shape, torch.float32))
comp.set_timing(True)
out = torch.empty(u.numel() * 4 + (1 << 20), dtype=torch.uint8, device="cuda")
res = {}
for mode in ["0", "1", "2", "0"]:
    os.environ["MGX_F3_DBG"] = mode
    ts = []
    for it in range(5):
        try:
            comp.Compress(u, mg.REL, tol, float("inf"), out=out)
        except Exception as e:  # noqa: BLE001
            pass
        ts.append(comp.elapsed_ms("c_norm", "c_refactor"))
    res.setdefault(mode, []).append(round(float(np.median(ts[1:])), 3))
print(json.dumps({"tol": tol, "refactor_quantize_ms by MGX_F3_DBG (0 normal, 1 no compute, 2 no loads)": res}))
