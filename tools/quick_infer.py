"""Scratch diagnostics on the GPU box: error distribution vs the oracle and a first timing."""
import sys, os, time
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as e
import importlib
pkg, orc = e.load_package(), e.load_oracle()
synth = importlib.import_module("nrc_b200.synthetic")
for enc in (0, 1):
    net = pkg.Network(); net.init(0, enc)
    master = net.get_params(pkg.PARAMS_MASTER_F32)
    if enc == 1:
        nm = net.n_params()[1]; master[nm:] = np.random.default_rng(5).uniform(-1, 1, master.size - nm).astype(np.float32)
    w16 = master.astype(np.float16); net.set_params(pkg.PARAMS_EMA_F16, w16)
    n = 20000
    q = synth.cornell_queries(n, seed=1)
    qd = torch.from_numpy(q).cuda(); od = torch.zeros(n, 3, device="cuda")
    net.infer(qd, od, n); torch.cuda.synchronize()
    got = od.cpu().numpy()
    spec = orc.make_spec(enc)
    r0 = orc.infer(spec, w16, q, 0); r1 = orc.infer(spec, w16, q, 1)
    for name, a, b in (("gpu vs oracle(fp32acc)", got, r0), ("oracle fp16acc vs fp32acc", r1, r0), ("gpu vs oracle(fp16acc)", got, r1)):
        d = np.abs(a - b)
        big = np.abs(b) > 0.02
        print(f"enc{enc} {name}: exact {np.mean(d == 0):.3f} abs max {d.max():.2e} p99.9 {np.quantile(d, .999):.2e} "
              f"rel(|ref|>0.02) max {np.max(d[big] / np.abs(b[big])):.2e}  mean|ref| {np.abs(b).mean():.3f}")
    # timing
    for N in (1 << 20, 3744000):
        q = synth.cornell_queries(4096, seed=2)
        qd = torch.from_numpy(np.tile(q, (N // 4096 + 1, 1))[:N].copy()).cuda(); od = torch.zeros(N, 3, device="cuda")
        for _ in range(3): net.infer(qd, od, N)
        torch.cuda.synchronize()
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
        ev[0].record()
        for _ in range(10): net.infer(qd, od, N)
        ev[1].record(); torch.cuda.synchronize()
        ms = ev[0].elapsed_time(ev[1]) / 10
        fl = 2 * (64 * (80 if enc == 0 else 64) + 4 * 4096 + 1024)
        print(f"enc{enc} N={N}: {ms*1e3:.1f} us  {N/ms/1e6:.2f} Gq/s  {N*fl/ms/1e9:.0f} TFLOP/s  {N*72/ms/1e6:.0f} GB/s")
    net.destroy()
