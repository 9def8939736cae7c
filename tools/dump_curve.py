"""Runs the 1000-step golden training schedule on the CUDA path and saves the losses (gpurun_out/curve_*.npy)."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
import __graft_entry__ as entry
import golden_util as gu
pkg = entry.load_package()
for enc, name in ((0, "freq"), (1, "hash")):
    net = pkg.Network(); net.init(0, enc)
    net.load_master(gu.seeded_master(enc, net.n_params()[0], False))
    got = np.zeros(gu.CURVE_STEPS, dtype=np.float32)
    for s in range(gu.CURVE_STEPS):
        q, t = gu.batch(s)
        got[s] = net.train(torch.from_numpy(q).cuda(), torch.from_numpy(t).cuda(), want_loss=True)
    np.save(os.path.join(ROOT, "gpurun_out", f"curve_{name}.npy"), got)
    net.destroy()
