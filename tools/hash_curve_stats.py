"""Run-to-run spread of OUR Hash loss curve (fp32 atomics make it non-deterministic, like the reference's fp16 atomics):
repeats the 1000-step golden curve N times and prints the statistics tests/test_golden.py gates on."""
import os, sys, json
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests")); sys.path.insert(0, ROOT)
import golden_util as gu
pkg = gu.entry.load_package()
g = gu.load("ref_curve_hash.npz")
ref, rr = g["losses"], g["losses_rerun"]
batches = [tuple(torch.from_numpy(a).cuda() for a in gu.batch(s)) for s in range(gu.CURVE_STEPS)]
rw, rrw = ref.reshape(-1, 50).mean(1), rr.reshape(-1, 50).mean(1)
rms = lambda a, b: float(np.sqrt(np.mean((np.log(a[3:]) - np.log(b[3:])) ** 2)))
rows, curves = [], []
for r in range(int(sys.argv[1]) if len(sys.argv) > 1 else 8):
    net = pkg.Network(); net.init(0, 1)
    net.load_master(gu.seeded_master(1, net.n_params()[0], False))
    got = np.array([net.train(q, t, want_loss=True) for q, t in batches], dtype=np.float32)
    net.destroy()
    gw = got.reshape(-1, 50).mean(1)
    dev = np.abs(gw / rw - 1)
    bias = abs(float(np.log(got[150:]).mean() - np.log(ref[150:]).mean()))
    med = np.abs(np.median(got.reshape(-1, 50), 1) / np.median(ref.reshape(-1, 50), 1) - 1)
    rows.append({"dev3": dev[:3].tolist(), "bias": bias, "rms": rms(gw, rw), "med_dev_max": float(med.max()), "med3": med[:3].tolist()})
    curves.append(got)
    print(json.dumps(rows[-1]))
sd = np.abs(rrw / rw - 1)
medr = np.abs(np.median(rr.reshape(-1, 50), 1) / np.median(ref.reshape(-1, 50), 1) - 1)
print("reference rerun:", json.dumps({"dev3": sd[:3].tolist(), "bias": abs(float(np.log(rr[150:]).mean() - np.log(ref[150:]).mean())),
                                      "rms": rms(rrw, rw), "med_dev_max": float(medr.max()), "med3": medr[:3].tolist()}))
c = np.stack(curves)
cw = c.reshape(c.shape[0], -1, 50).mean(2)
print("ours vs ours (run 0 as the reference): dev3", [np.abs(cw[i] / cw[0] - 1)[:3].round(4).tolist() for i in range(1, c.shape[0])])
np.savez_compressed(os.path.join(ROOT, "gpurun_out", "hash_curves_ours.npz"), curves=c)
