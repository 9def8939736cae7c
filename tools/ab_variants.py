"""A/B of library builds on the GPU box: parity against the oracle + CUDA-event timing of the query kernel and of the frame's
training launch, one subprocess per build (NRC_B200_LIB selects the .so).

    make -C neural-radiance-caching_b200/csrc variants        (here: cross-compiles build/variants/libnrc_b200_v{0,1,2}.so)
    gpurun -- python tools/ab_variants.py                     (default build + every build/variants/*.so)
"""
import glob
import importlib
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def child():
    import numpy as np
    import torch
    import __graft_entry__ as e
    pkg, orc = e.load_package(), e.load_oracle()
    synth = importlib.import_module("nrc_b200.synthetic")
    res = {"lib": os.path.basename(pkg.LIB_PATH)}
    frame = synth.cornell_frame(2560, 1440, n_tail=57600)
    n = frame.shape[0]
    q_d = torch.from_numpy(frame).cuda()
    out_d = torch.empty((n, 3), device="cuda")
    tq = synth.cornell_queries(4 * pkg.BATCH_SIZE, seed=4)
    tt = synth.radiance_targets(tq, seed=4)
    tq_d, tt_d = torch.from_numpy(tq).cuda(), torch.from_numpy(tt).cuda()
    rng = np.random.default_rng(17)
    wide = np.empty((8192, 15), dtype=np.float32)
    wide[:, 0:3] = rng.uniform(-3.0, 4.0, (8192, 3)); wide[:, 3:7] = rng.uniform(-12.0, 12.0, (8192, 4))
    wide[:, 7:9] = rng.uniform(-2.0, 3.0, (8192, 2)); wide[:, 9:15] = rng.uniform(0.0, 8.0, (8192, 6))
    for enc in (0, 1):
        net = pkg.Network()
        net.init(0, enc)
        master = net.get_params(pkg.PARAMS_MASTER_F32)
        if enc == 1:
            nm = net.n_params()[1]
            master[nm:] = np.random.default_rng(5).uniform(-1, 1, master.size - nm).astype(np.float32)
        w16 = master.astype(np.float16)
        net.set_params(pkg.PARAMS_EMA_F16, w16)
        spec = orc.make_spec(enc)
        for name, q in (("cornell", synth.cornell_queries(20000, seed=1)), ("wide", wide)):
            qd = torch.from_numpy(q).cuda()
            od = torch.zeros(q.shape[0], 3, device="cuda")
            net.infer(qd, od, q.shape[0])
            torch.cuda.synchronize()
            got, ref = od.cpu().numpy(), orc.infer(spec, w16 if name == "cornell" else w16, q, 0)
            err = np.abs(got - ref) / np.maximum(np.abs(ref), 0.05)
            res[f"enc{enc}_{name}_max_rel"] = float(err.max())
            res[f"enc{enc}_{name}_exact"] = float(np.mean(got == ref))
        for _ in range(5):
            net.infer(q_d, out_d, n)
        torch.cuda.synchronize()
        best = []
        for _ in range(3):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            for _ in range(20):
                net.infer(q_d, out_d, n)
            b.record()
            torch.cuda.synchronize()
            best.append(a.elapsed_time(b) / 20)
        res[f"enc{enc}_infer_us"] = min(best) * 1e3
        # incoherent batch (independent random surface points, 2^21 queries = 126 MB > L2 share): the worst case for the gathers
        qr = torch.from_numpy(np.concatenate([synth.cornell_queries(1 << 20, seed=s_) for s_ in (1, 2)])).cuda()
        for _ in range(3):
            net.infer(qr, out_d, 1 << 21)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(10):
            net.infer(qr, out_d, 1 << 21)
        b.record()
        torch.cuda.synchronize()
        res[f"enc{enc}_infer_random_2M_us"] = a.elapsed_time(b) / 10 * 1e3
        del qr
        for _ in range(3):
            net.train_batches(tq_d.data_ptr(), tt_d.data_ptr(), 4)
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(20):
            net.train_batches(tq_d.data_ptr(), tt_d.data_ptr(), 4)
        b.record()
        torch.cuda.synchronize()
        res[f"enc{enc}_train4_us"] = a.elapsed_time(b) / 20 * 1e3
        net.destroy()
    print("AB " + json.dumps(res))


def main():
    if os.environ.get("NRC_AB_CHILD"):
        return child()
    libs = [None] + sorted(glob.glob(os.path.join(ROOT, "build", "variants", "*.so")))
    rows = []
    for lib in libs:
        env = dict(os.environ, NRC_AB_CHILD="1")
        if lib:
            env["NRC_B200_LIB"] = lib
        r = subprocess.run([sys.executable, os.path.abspath(__file__)], env=env, capture_output=True, text=True, timeout=600)
        line = [l for l in r.stdout.splitlines() if l.startswith("AB ")]
        if not line:
            print(f"{lib}: FAILED\n{r.stdout[-2000:]}\n{r.stderr[-2000:]}")
            continue
        rows.append(json.loads(line[0][3:]))
        print(json.dumps(rows[-1]))
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    json.dump(rows, open(os.path.join(ROOT, "gpurun_out", "ab_variants.json"), "w"), indent=1)


if __name__ == "__main__":
    main()
