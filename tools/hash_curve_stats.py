"""Run-to-run spread of OUR Hash loss curve (atomics make it non-deterministic, like the reference's fp16 atomics): repeats the
1000-step golden curve N times in each arithmetic and prints, per run, exactly the quantities tests/test_golden.py gates on
(test_cuda_loss_curve_1000_steps_matches_reference), so that the gates can be set from the measured distribution.
    python tools/hash_curve_stats.py [N runs per arithmetic, default 10]"""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, ROOT)
import golden_util as gu  # noqa: E402

pkg = gu.entry.load_package()
g = gu.load("ref_curve_hash.npz")
ref, rr = g["losses"], g["losses_rerun"]
runs = g["losses_runs"].reshape(g["losses_runs"].shape[0], -1, 50).mean(2)
mean, sd = runs.mean(0), runs.std(0, ddof=1)
batches = [tuple(torch.from_numpy(a).cuda() for a in gu.batch(s)) for s in range(gu.CURVE_STEPS)]
wmed = lambda a: np.median(a.reshape(-1, 50), axis=1)
rms = lambda a, b: float(np.sqrt(np.mean((np.log(wmed(a)[3:]) - np.log(wmed(b)[3:])) ** 2)))
q_inf = torch.from_numpy(gu.infer_queries()).cuda()
h = gu.N_INFER // 2
rel_err = lambda o, r: np.abs(o - r) / np.maximum(np.abs(r), 1e-3)
n_runs = int(sys.argv[1]) if len(sys.argv) > 1 else 10
print("reference over", runs.shape[0], "runs: sd/mean windows 0..2", (sd / mean)[:3].round(4).tolist(),
      "| rerun: tail bias", round(abs(float(np.log(rr[150:]).mean() - np.log(ref[150:]).mean())), 4), "rms", round(rms(rr, ref), 4))
out = {}
for ref_arith in (True, False):
    rows = []
    for r in range(n_runs):
        net = pkg.Network()
        net.init(0, 1)
        if ref_arith:
            net.set_accumulator(pkg.ACC_FP16)
        net.load_master(gu.seeded_master(1, net.n_params()[0], False))
        got = np.array([net.train(q, t, want_loss=True) for q, t in batches], dtype=np.float32)
        o = torch.zeros((q_inf.shape[0], 3), device="cuda")
        net.infer(q_inf, o, q_inf.shape[0])
        torch.cuda.synchronize()
        net.destroy()
        gw = got.reshape(-1, 50).mean(1)
        rel = gw[:3] / mean[:3] - 1
        oo, rad = o.cpu().numpy()[:h], g["radiance"][:h]
        rows.append({"dev0": float(abs(gw[0] / ref.reshape(-1, 50).mean(1)[0] - 1)), "rel": rel.round(4).tolist(),
                     "z": ((gw[:3] - mean[:3]) / sd[:3]).round(2).tolist(),
                     "bias": round(abs(float(np.log(got[150:]).mean() - np.log(ref[150:]).mean())), 4), "rms": round(rms(got, ref), 4),
                     "corr": round(float(np.corrcoef(oo.ravel(), rad.ravel())[0, 1]), 5), "median_rel": round(float(np.median(rel_err(oo, rad))), 4)})
        print("reference arithmetic" if ref_arith else "default arithmetic  ", json.dumps(rows[-1]))
    out["ref_arith" if ref_arith else "default"] = rows
    for k in ("bias", "rms", "corr", "median_rel"):
        v = np.array([x[k] for x in rows])
        print(f"   {k}: min {v.min():.4f} max {v.max():.4f}")
    rel = np.array([x["rel"] for x in rows])
    print("   rel windows 0..2: min", rel.min(0).round(4).tolist(), "max", rel.max(0).round(4).tolist())
json.dump(out, open(os.path.join(ROOT, "gpurun_out", "hash_curve_stats.json"), "w"), indent=1)
