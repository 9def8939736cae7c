"""Smallest program that launches the training step kernel (for `ncu -k regex:nrc_train_step -s 6 -c 1`):
    python tools/ncu_train.py <0|1 encoding> [single|multi]       (NRC_B200_LIB selects an A/B build)"""
import importlib
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as e  # noqa: E402


def main():
    import torch
    enc = int(sys.argv[1]) if len(sys.argv) > 1 else 0
    multi = len(sys.argv) > 2 and sys.argv[2] == "multi"
    pkg = e.load_package()
    synth = importlib.import_module("nrc_b200.synthetic")
    tq = synth.cornell_queries(4 * pkg.BATCH_SIZE, seed=4)
    tt = synth.radiance_targets(tq, seed=4)
    tq_d, tt_d = torch.from_numpy(tq).cuda(), torch.from_numpy(tt).cuda()
    net = pkg.Network()
    net.init(0, enc)
    for _ in range(3):
        if multi:
            net.train_batches(tq_d.data_ptr(), tt_d.data_ptr(), 4)
        else:
            for b in range(4):
                net.train(tq_d[b * pkg.BATCH_SIZE:], tt_d[b * pkg.BATCH_SIZE:])
    torch.cuda.synchronize()
    net.destroy()
    print("ok")


if __name__ == "__main__":
    main()
