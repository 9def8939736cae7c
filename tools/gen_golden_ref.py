"""Generates tests/golden/ref_*.npz by running the UNMODIFIED reference (nrc::Network + tiny-cuda-nn, compiled by
oracle/ref/Makefile.ref into oracle/_ref/libnrc_ref.so) on a B200:

    gpurun -- 'python tools/gen_golden_ref.py gpurun_out/golden'   then copy gpurun_out/golden/*.npz to tests/golden/

Everything that can be regenerated from a seed (weights, queries, targets) is NOT stored; see tests/golden_util.py
for the generators the tests and this script share.  Stored: the reference's outputs.
  ref_infer_{freq,hash}.npz : radiance of the reference for N_INFER seeded queries with seeded weights.  The weights
        reach the reference's inference (EMA) buffer through one training step at learning rate 0: weights stay
        put and the EMA after step 1 equals the fp16 weights exactly (ema.h:107-108).
  ref_step1_{freq,hash}.npz : loss and fp16 gradient vector (Freq: all; Hash: MLP part + a strided sample of the
        grid part) of the first training step from the seeded weights.
  ref_curve_{freq,hash}.npz : per-step losses of 1000 training steps on seeded batches, the reference's final
        fp16 EMA weights (Freq) and its radiance for the seeded queries after training.
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle", "ref"))
sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    import torch
    import ref_gpu
    import golden_util as gu
    out_dir = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "gpurun_out", "golden")
    n_steps = int(sys.argv[2]) if len(sys.argv) > 2 else gu.CURVE_STEPS
    os.makedirs(out_dir, exist_ok=True)
    torch.cuda.set_device(0)
    for enc, name in ((0, "freq"), (1, "hash")):
        # ---------------- inference with seeded weights
        net = ref_gpu.RefNetwork(enc, 0)
        master = gu.seeded_master(enc, net.n, visible_grid=True)
        net.set_master(master)
        net.set_lr(0.0)
        q0, t0 = gu.batch(0)
        q0_d, t0_d = torch.from_numpy(q0).cuda(), torch.from_numpy(t0).cuda()
        net.train(q0_d.data_ptr(), t0_d.data_ptr(), want_loss=True)
        ema = net.get_params(2)
        assert np.array_equal(ema, master.astype(np.float16)), "EMA after an lr=0 step must equal the fp16 weights"
        q = gu.infer_queries()
        q_d = torch.from_numpy(q).cuda()
        o_d = torch.zeros((q.shape[0], 3), device="cuda")
        net.infer(q_d.data_ptr(), o_d.data_ptr(), q.shape[0])
        torch.cuda.synchronize()
        np.savez_compressed(os.path.join(out_dir, f"ref_infer_{name}.npz"), radiance=o_d.cpu().numpy())
        net.destroy()

        # ---------------- first training step: loss + gradients
        net = ref_gpu.RefNetwork(enc, 0)
        master = gu.seeded_master(enc, net.n, visible_grid=True)
        net.set_master(master)
        loss1 = net.train(q0_d.data_ptr(), t0_d.data_ptr(), want_loss=True)
        g = net.get_params(3)
        n_mlp = gu.n_mlp(enc)
        np.savez_compressed(os.path.join(out_dir, f"ref_step1_{name}.npz"), loss=np.float32(loss1), grad_mlp=g[:n_mlp],
                            grad_grid_sample=g[n_mlp::gu.GRID_SAMPLE_STRIDE],
                            grad_grid_nonzero=np.int64((g[n_mlp:] != 0).sum()),
                            grad_grid_abs_sum=np.float64(np.abs(g[n_mlp:].astype(np.float64)).sum()))
        net.destroy()

        # ---------------- loss curve over n_steps from the seeded init (grid at its U(-1e-4,1e-4)-like init).
        # Run twice: the Hash config scatters gradients with fp16 atomics (grid.h:301-319), so the reference is
        # not reproducible run to run; `losses_rerun` lets the tests scale their tolerance by the reference's own spread.
        curves = []
        for rerun in range(2):
            net = ref_gpu.RefNetwork(enc, 0)
            master = gu.seeded_master(enc, net.n, visible_grid=False)
            net.set_master(master)
            losses = np.zeros(n_steps, dtype=np.float32)
            for s in range(n_steps):
                qs, ts = gu.batch(s)
                qs_d, ts_d = torch.from_numpy(qs).cuda(), torch.from_numpy(ts).cuda()
                losses[s] = net.train(qs_d.data_ptr(), ts_d.data_ptr(), want_loss=True)
            curves.append(losses)
            if rerun == 0:
                net.infer(q_d.data_ptr(), o_d.data_ptr(), q.shape[0])
                torch.cuda.synchronize()
                radiance = o_d.cpu().numpy()
                extra = {"ema": net.get_params(2)} if enc == 0 else {}
            net.destroy()
        losses = curves[0]
        np.savez_compressed(os.path.join(out_dir, f"ref_curve_{name}.npz"), losses=curves[0], losses_rerun=curves[1],
                            radiance=radiance, **extra)
        print(name, "loss1", loss1, "curve", losses[:3], losses[-3:])


if __name__ == "__main__":
    main()
