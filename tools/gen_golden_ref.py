"""Generates tests/golden/ref_*.npz by running the UNMODIFIED reference (nrc::Network + tiny-cuda-nn and the frame helper
kernels of nrc/shaders/nrc_helpers.cu, compiled by oracle/ref/Makefile.ref into oracle/_ref/libnrc_ref.so) on a B200:

    gpurun -- 'python tools/gen_golden_ref.py gpurun_out/golden [steps] [reruns] [sections]'
    then copy gpurun_out/golden/*.npz to tests/golden/

Everything that can be regenerated from a seed (weights, queries, targets, chains) is NOT stored; see tests/golden_util.py
for the generators the tests and this script share.  Stored: the reference's outputs.
  ref_infer_{freq,hash}.npz : radiance of the reference for N_INFER seeded queries with seeded weights.  The weights
        reach the reference's inference (EMA) buffer through one training step at learning rate 0: weights stay
        put and the EMA after step 1 equals the fp16 weights exactly (ema.h:107-108).
  ref_step1_{freq,hash}.npz : loss and fp16 gradient vector (Freq: all; Hash: MLP part + a strided sample of the
        grid part) of the first training step from the seeded weights.
  ref_curve_{freq,hash}.npz : per-step losses of 1000 training steps on seeded batches, REPEATED `reruns` times (the Hash
        config scatters gradients with fp16 atomics and is not reproducible run to run: `losses_runs` is the reference's
        own distribution), the reference's final fp16 EMA weights of run 0 (both encodings: the trained cache) and its
        radiance for the seeded queries after training.
  ref_helpers.npz : outputs of the reference's accumulate_render_radiance / propagate_train_radiance / permute_train_data
        kernels on the seeded inputs of golden_util.helper_inputs().
  ref_frame_{freq,hash}.npz : one small frame through the whole NRC sequence of Device::render (infer -> accumulate ->
        propagate -> permute -> 4 x train), reference network + reference helper kernels.
Sections (comma separated, default all): infer,step1,curve,helpers,frame
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
    n_runs = int(sys.argv[3]) if len(sys.argv) > 3 else 12
    sections = set((sys.argv[4] if len(sys.argv) > 4 else "infer,step1,curve,helpers,frame").split(","))
    os.makedirs(out_dir, exist_ok=True)
    torch.cuda.set_device(0)
    dev = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()

    batches = None
    if "curve" in sections:     # the training batches of the curve, generated once and kept on the device
        batches = [tuple(dev(x) for x in gu.batch(s)) for s in range(n_steps)]

    for enc, name in ((0, "freq"), (1, "hash")):
        q0, t0 = gu.batch(0)
        q0_d, t0_d = dev(q0), dev(t0)
        q = gu.infer_queries()
        q_d = dev(q)
        o_d = torch.zeros((q.shape[0], 3), device="cuda")
        if "infer" in sections:
            # ---------------- inference with seeded weights
            net = ref_gpu.RefNetwork(enc, 0)
            master = gu.seeded_master(enc, net.n, visible_grid=True)
            net.set_master(master)
            net.set_lr(0.0)
            net.train(q0_d.data_ptr(), t0_d.data_ptr(), want_loss=True)
            ema = net.get_params(2)
            assert np.array_equal(ema, master.astype(np.float16)), "EMA after an lr=0 step must equal the fp16 weights"
            net.infer(q_d.data_ptr(), o_d.data_ptr(), q.shape[0])
            torch.cuda.synchronize()
            np.savez_compressed(os.path.join(out_dir, f"ref_infer_{name}.npz"), radiance=o_d.cpu().numpy())
            net.destroy()

        if "step1" in sections:
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
            print(name, "loss1", loss1)

        if "curve" in sections:
            # ---------------- loss curve over n_steps from the seeded init (grid at its U(-1e-4,1e-4)-like init), n_runs
            # times: the reference's own run-to-run distribution (Hash: fp16 atomics, grid.h:301-319).
            curves = []
            for run in range(n_runs):
                net = ref_gpu.RefNetwork(enc, 0)
                master = gu.seeded_master(enc, net.n, visible_grid=False)
                net.set_master(master)
                losses = np.zeros(n_steps, dtype=np.float32)
                for s in range(n_steps):
                    losses[s] = net.train(batches[s][0].data_ptr(), batches[s][1].data_ptr(), want_loss=True)
                curves.append(losses)
                if run == 0:
                    net.infer(q_d.data_ptr(), o_d.data_ptr(), q.shape[0])
                    torch.cuda.synchronize()
                    radiance = o_d.cpu().numpy()
                    ema = net.get_params(2)
                net.destroy()
            runs = np.stack(curves)
            np.savez_compressed(os.path.join(out_dir, f"ref_curve_{name}.npz"), losses=runs[0], losses_rerun=runs[1],
                                losses_runs=runs, radiance=radiance, ema=ema)
            w = runs.reshape(n_runs, -1, 50).mean(2)
            print(name, "curve windows 0..3 min/max over runs:", w[:, :4].min(0), w[:, :4].max(0))

        if "frame" in sections:
            # ---------------- one frame through the reference's whole NRC sequence (Device.cpp:2495-2515)
            f = gu.frame_inputs(enc)
            net = ref_gpu.RefNetwork(enc, 0)
            net.set_master(gu.seeded_master(enc, net.n, visible_grid=True))
            lr = 1e-3 if enc == 0 else 1e-2
            net.set_lr(0.0)                                      # weights -> EMA buffer (see above); Adam moments move
            net.train(q0_d.data_ptr(), t0_d.data_ptr(), want_loss=True)
            net.set_lr(lr)
            n_px, (tx, ty) = f["n_px"], f["tiles"]
            n_q = f["queries"].shape[0]
            assert n_q % 256 == 0
            fq_d, res_d = dev(f["queries"]), torch.zeros((n_q, 3), device="cuda")
            thr_d, out_d = dev(f["throughput"]), dev(f["output"])
            ev_d, rec_d = dev(f["ev"]), dev(f["rec"])
            tq_d, tt_d = dev(f["train_q"]), dev(f["train_t"])
            tq2_d, tt2_d = torch.zeros_like(tq_d), torch.zeros_like(tt_d)
            perm_d = dev(f["perm"])
            net.infer(fq_d.data_ptr(), res_d.data_ptr(), n_q)
            torch.cuda.synchronize()
            ref_gpu.accumulate_render_radiance(res_d.data_ptr(), thr_d.data_ptr(), out_d.data_ptr(), gu.FRAME_W, gu.FRAME_H, 0, 2)
            ref_gpu.propagate_train_radiance(ev_d.data_ptr(), res_d.data_ptr() + n_px * 12, rec_d.data_ptr(), tt_d.data_ptr(), tx, ty)
            ref_gpu.permute_train_data(tq_d.data_ptr(), tq2_d.data_ptr(), tt_d.data_ptr(), tt2_d.data_ptr(), perm_d.data_ptr(), f["used"])
            losses = [net.train(tq2_d.data_ptr() + b * 16384 * 60, tt2_d.data_ptr() + b * 16384 * 12, want_loss=True) for b in range(4)]
            np.savez_compressed(os.path.join(out_dir, f"ref_frame_{name}.npz"), results=res_d.cpu().numpy()[::7],
                                output=out_d.cpu().numpy()[::5], targets=tt_d.cpu().numpy()[:f["used"]][::3],
                                permuted_q_digest=gu.digest(tq2_d.cpu().numpy()), permuted_t_digest=gu.digest(tt2_d.cpu().numpy()),
                                losses=np.array(losses, dtype=np.float32), ema_mlp=net.get_params(2)[:gu.n_mlp(enc)])
            net.destroy()
            print(name, "frame losses", losses)

    if "helpers" in sections:
        # ---------------- the reference's frame helper kernels on seeded inputs
        h = gu.helper_inputs()
        out = {}
        rad_d, thr_d = dev(h["acc_radiance"]), dev(h["acc_throughput"])
        for mode, it in gu.ACC_CASES:
            o_d = dev(h["acc_output"])
            ref_gpu.accumulate_render_radiance(rad_d.data_ptr(), thr_d.data_ptr(), o_d.data_ptr(), gu.HELPER_W, gu.HELPER_H, mode, it)
            out[f"acc_m{mode}_i{it}"] = o_d.cpu().numpy()
        tg_d = dev(h["prop_targets"])
        ref_gpu.propagate_train_radiance(dev(h["prop_ev"]).data_ptr(), dev(h["prop_end_radiance"]).data_ptr(), dev(h["prop_rec"]).data_ptr(),
                                         tg_d.data_ptr(), gu.HELPER_TILES_X, gu.HELPER_TILES_Y)
        out["prop_targets"] = tg_d.cpu().numpy()
        pq_d, pt_d, perm_d = dev(h["perm_q"]), dev(h["perm_t"]), dev(h["perm"])
        for n_rec in gu.PERMUTE_CASES:
            dq_d, dt_d = torch.zeros_like(pq_d), torch.zeros_like(pt_d)
            ref_gpu.permute_train_data(pq_d.data_ptr(), dq_d.data_ptr(), pt_d.data_ptr(), dt_d.data_ptr(), perm_d.data_ptr(), n_rec)
            dq, dt = dq_d.cpu().numpy(), dt_d.cpu().numpy()
            out[f"perm_{n_rec}_q_digest"], out[f"perm_{n_rec}_t_digest"] = gu.digest(dq), gu.digest(dt)
            out[f"perm_{n_rec}_q_sample"], out[f"perm_{n_rec}_t_sample"] = dq[::97], dt[::97]
        np.savez_compressed(os.path.join(out_dir, "ref_helpers.npz"), **out)
        print("helpers ok:", sorted(out)[:4], "...")


if __name__ == "__main__":
    main()
