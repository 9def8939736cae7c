"""Smallest program that launches the query kernel on the 2K frame (for `ncu -k regex:nrc_infer -s 3 -c 1`):
    python tools/ncu_infer.py <0|1 encoding> [frame|random]       (NRC_B200_LIB selects an A/B build)"""
import importlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as e  # noqa: E402


def main():
    import torch
    enc = int(sys.argv[1]) if len(sys.argv) > 1 else 0
    order = sys.argv[2] if len(sys.argv) > 2 else "frame"
    pkg = e.load_package()
    synth = importlib.import_module("nrc_b200.synthetic")
    q = synth.cornell_frame(2560, 1440, n_tail=57600) if order == "frame" else np.concatenate(
        [synth.cornell_queries(1 << 20, seed=s) for s in (1, 2)])
    n = q.shape[0]
    q_d, out_d = torch.from_numpy(q).cuda(), torch.empty((n, 3), device="cuda")
    net = pkg.Network()
    net.init(0, enc)
    master = net.get_params(pkg.PARAMS_MASTER_F32)
    if enc == 1:
        nm = net.n_params()[1]
        master[nm:] = np.random.default_rng(5).uniform(-1, 1, master.size - nm).astype(np.float32)
    net.set_params(pkg.PARAMS_EMA_F16, master.astype(np.float16))
    for _ in range(5):
        net.infer(q_d, out_d, n)
    torch.cuda.synchronize()
    net.destroy()
    print("ok", n)


if __name__ == "__main__":
    main()
