"""Prints the phase timeline (clock64 deltas, median over tiles) of the fused training step kernel."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import __graft_entry__ as entry
pkg = entry.load_package(); synth = entry.load_package_synthetic()
for enc in (0, 1):
    net = pkg.Network(); net.init(0, enc)
    # four different batches in turn, like a frame: an entry of the hash grid that a batch misses keeps a lagging step count,
    # which is the steady state of the optimizer phase (with ONE repeated batch every counter equals the global step)
    q = synth.cornell_queries(4 * 16384, seed=1); t = synth.radiance_targets(q, seed=1)
    qa, ta = torch.from_numpy(q).cuda(), torch.from_numpy(t).cuda()
    qs, ts = [qa[b * 16384:(b + 1) * 16384] for b in range(4)], [ta[b * 16384:(b + 1) * 16384] for b in range(4)]
    qd, td = qs[0], ts[0]
    for i in range(24): net.train(qs[i % 4], ts[i % 4])
    net.debug_train_timeline(True)
    for i in range(3): net.train(qs[i % 4], ts[i % 4])
    tl = net.debug_train_timeline(True, read=True).astype(np.float64)
    d = np.diff(tl[:, :7], axis=1) / 1.965e3   # us at 1965 MHz
    names = ["setup+encode", "forward", "backward+wgrad", "scatter+dump", "grid barrier", "reduce+adam"]
    print("enc", enc, " ".join(f"{n}={np.median(d[:, i]):.2f}us(max {d[:, i].max():.2f})" for i, n in enumerate(names)),
          f"total={np.median(tl[:, 6] - tl[:, 0]) / 1.965e3:.2f}us",
          f"half1-scatter-done at +{np.median(tl[:, 7] - tl[:, 3]) / 1.965e3:.2f}us after backward")
    # the frame's four steps in ONE launch: the stamps are those of the launch's second step (slot 0 = top of that step)
    for _ in range(3): net.train_batches(qa.data_ptr(), ta.data_ptr(), 4)
    tl = net.debug_train_timeline(True, read=True).astype(np.float64)
    d = np.diff(tl[:, :7], axis=1) / 1.965e3
    names_m = ["(barrier 2 +) weights + encode", "forward", "backward+wgrad", "scatter+dump", "grid barrier", "reduce+adam"]
    print("enc", enc, "4-step launch, step 2:", " ".join(f"{n}={np.median(d[:, i]):.2f}us(max {d[:, i].max():.2f})" for i, n in enumerate(names_m)),
          f"total={np.median(tl[:, 6] - tl[:, 0]) / 1.965e3:.2f}us")
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    net.debug_train_timeline(False)
    torch.cuda.synchronize(); e0.record()
    for i in range(200): net.train(qs[i % 4], ts[i % 4])
    e1.record(); torch.cuda.synchronize()
    print("   back-to-back step:", e0.elapsed_time(e1) / 200 * 1e3, "us")
    net.destroy()
