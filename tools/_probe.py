import os, sys, numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as entry
pkg = entry.load_package(); synth = entry.load_package_synthetic(); orc = entry.load_oracle()
for enc in (0, 1):
    net = pkg.Network(); net.init(0, enc, seed=1337)
    master = net.get_params(pkg.PARAMS_MASTER_F32)
    if enc == 1:
        nm = net.n_params()[1]
        master[nm:] = np.random.default_rng(5).uniform(-1, 1, master.size - nm).astype(np.float32)
    w16 = master.astype(np.float16)
    net.set_params(pkg.PARAMS_EMA_F16, w16)
    q = synth.cornell_queries(20000, seed=3)
    qd = torch.from_numpy(q).cuda(); od = torch.empty(20000, 3, device="cuda")
    net.infer(qd, od, 20000); torch.cuda.synchronize()
    got = od.cpu().numpy()
    r32 = orc.infer(orc.make_spec(enc), w16, q, acc_mode=0); r16 = orc.infer(orc.make_spec(enc), w16, q, acc_mode=1)
    f = lambda a, b: (np.abs(a - b).max(), float(np.mean(a == b)), (np.abs(a - b) / np.maximum(np.abs(b), 0.05)).max())
    print("enc", enc, "vs fp32-acc oracle", f(got, r32), "vs fp16-acc oracle", f(got, r16), "oracle16 vs oracle32", f(r16, r32), "mean|ref|", np.abs(r32).mean())
