import os, sys, numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as entry
pkg = entry.load_package(); synth = entry.load_package_synthetic(); orc = entry.load_oracle()
for log2 in (14, 15, 18, 21):
  for kind in ("uniform", "cornell"):
    net = pkg.Network(); net.init(0, 1, log2_hashmap_size=log2, seed=1337)
    master = net.get_params(pkg.PARAMS_MASTER_F32)
    nm = net.n_params()[1]
    master[nm:] = np.random.default_rng(5).uniform(-1, 1, master.size - nm).astype(np.float32)
    w16 = master.astype(np.float16)
    net.set_params(pkg.PARAMS_EMA_F16, w16)
    q = synth.uniform_queries(4096, seed=log2) if kind == "uniform" else synth.cornell_queries(4096, seed=log2)
    qd = torch.from_numpy(q).cuda(); od = torch.empty(4096, 3, device="cuda")
    net.infer(qd, od, 4096); torch.cuda.synchronize()
    got = od.cpu().numpy()
    spec = orc.make_spec(1, log2)
    r32 = orc.infer(spec, w16, q, acc_mode=0); r16 = orc.infer(spec, w16, q, acc_mode=1)
    print(log2, kind, "max|got-r32|", np.abs(got - r32).max(), "equal", np.mean(got == r32), "max|r16-r32|", np.abs(r16 - r32).max(), "mean|ref|", np.abs(r32).mean(), "n_params", net.n_params())
    net.destroy()
