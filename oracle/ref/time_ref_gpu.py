"""Times the reference's own tcnn build (oracle/_ref/libnrc_ref.so) on this GPU for one NRC update:
infer(nq) + 4 x train(16384), CUDA events, default stream.  Prints one JSON object.  Run as a subprocess by
`bench.py --ref-gpu` so that a crash of the legacy kernels cannot take the benchmark down."""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))


def main():
    import torch
    import ref_gpu
    import __graft_entry__ as entry
    enc, nq = int(sys.argv[1]), int(sys.argv[2])
    steps = int(sys.argv[3]) if len(sys.argv) > 3 else 20
    synth = entry.load_package_synthetic()
    torch.cuda.set_device(0)
    n_pad = (nq + 255) // 256 * 256
    q = torch.zeros((n_pad, 15), device="cuda")
    q[:min(nq, 1 << 20)] = torch.from_numpy(synth.cornell_queries(min(nq, 1 << 20), seed=1337)).cuda()
    if nq > (1 << 20):   # tile the first block
        reps = (nq + (1 << 20) - 1) // (1 << 20)
        q[:nq] = q[:1 << 20].repeat(reps, 1)[:nq]
    out = torch.zeros((n_pad, 3), device="cuda")
    tq_np = synth.cornell_queries(4 * 16384, seed=4242)
    tq = torch.from_numpy(tq_np).cuda()
    tt = torch.from_numpy(synth.radiance_targets(tq_np, seed=4242)).cuda()
    net = ref_gpu.RefNetwork(enc, 0)

    def update(loss):
        net.infer(q.data_ptr(), out.data_ptr(), nq)
        for b in range(4):
            net.train(tq.data_ptr() + b * 16384 * 60, tt.data_ptr() + b * 16384 * 12, want_loss=loss)

    res = {}
    for label, loss in (("async", False), ("loss_each_step", True)):   # the reference app reads the loss every step
        for _ in range(5):
            update(loss)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            update(loss)
        e1.record()
        torch.cuda.synchronize()
        res["ms_per_update_" + label] = e0.elapsed_time(e1) / steps
    # inference alone
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        net.infer(q.data_ptr(), out.data_ptr(), nq)
    e1.record()
    torch.cuda.synchronize()
    res["ms_infer"] = e0.elapsed_time(e1) / steps
    res["queries_per_s"] = nq / (res["ms_infer"] * 1e-3)
    res["samples_per_s"] = (nq + 65536) / (res["ms_per_update_async"] * 1e-3)
    res["what"] = "reference nrc::Network + tiny-cuda-nn compiled for sm_100 (mma.sync HMMA), same box, CUDA events"
    net.destroy()
    print(json.dumps(res))


if __name__ == "__main__":
    main()
