"""NEXT rows: accumulate / propagate / shuffle+permute / frame driver.
CPU part: the oracle restatements against hand-derived facts and the ABI's host-side pieces.
GPU part (@gpu): the CUDA kernels against the oracle -- bit-exact (byte moves and single fused multiply-adds),
the fused shuffle against permute-then-train, and nrc_b200_update against the explicit call sequence."""
import numpy as np
import pytest


def _chains(orc, n_tiles, n_records, rng, max_len=6):
    """Synthetic TrainingRecord chains: every tile's suffix end points at a chain of 0..max_len records."""
    rec = np.zeros(n_records, dtype=orc.RECORD_DTYPE)
    rec["propTo"] = -1
    ev = np.zeros(n_tiles, dtype=orc.END_VERTEX_DTYPE)
    ev["startTrainRecord"] = -1
    nxt = 0
    for t in range(n_tiles):
        length = int(rng.integers(0, max_len + 1))
        if nxt + length > n_records:
            break
        if length:
            ids = np.arange(nxt, nxt + length)
            rec["propTo"][ids[:-1]] = ids[1:]
            rec["propTo"][ids[-1]] = -1
            ev["startTrainRecord"][t] = ids[0]
            nxt += length
        ev["radianceMask"][t] = float(rng.integers(0, 2))
        ev["tileIndex"][t] = t
    rec["localThroughput"] = rng.uniform(0.0, 1.2, (n_records, 3)).astype(np.float32)
    return ev, rec, nxt


# ------------------------------------------------------------------ CPU
def test_permutation_is_a_bijection_and_matches_abi(orc, pkg):
    for seed in (0, 1, 12345, 2**32 - 1):
        perm = np.array([orc.permutation(i, seed) for i in range(65536)])
        assert np.array_equal(np.sort(perm), np.arange(65536))
        idx = np.arange(0, 65536, 257)
        assert all(pkg.permutation(int(i), seed) == perm[i] for i in idx)        # host side of the C ABI
    a = np.array([orc.permutation(i, 1) for i in range(4096)])
    b = np.array([orc.permutation(i, 2) for i in range(4096)])
    assert np.mean(a == b) < 0.01                                                  # the seed matters
    assert abs(np.corrcoef(np.arange(4096), a)[0, 1]) < 0.1                        # neighbours are decorrelated


def test_permute_duplicates_when_undersampled(orc, synth):
    q = synth.cornell_queries(1000, seed=1)
    t = synth.radiance_targets(q, seed=1)
    qd, td = orc.permute(q, t, 1000, 5)
    src = np.array([orc.permutation(d, 5) % 1000 for d in range(65536)])
    assert np.array_equal(qd, q[src]) and np.array_equal(td, t[src])
    counts = np.bincount(src, minlength=1000)
    assert counts.min() >= 50 and counts.max() <= 80                              # 65.5 copies of each record on average
    qd0, _ = orc.permute(q, t, 0, 5)
    assert (qd0 == 0).all()                                                        # numRecords <= 0: untouched (nrc_helpers.cu:238)


def test_propagate_closed_form(orc):
    rec = np.zeros(3, dtype=orc.RECORD_DTYPE)
    rec["propTo"] = [1, 2, -1]
    rec["localThroughput"] = [[0.5] * 3, [2.0] * 3, [0.25] * 3]
    ev = np.zeros(2, dtype=orc.END_VERTEX_DTYPE)
    ev["startTrainRecord"] = [0, -1]
    ev["radianceMask"] = [1.0, 1.0]
    tgt = np.array([[1, 1, 1], [10, 10, 10], [100, 100, 100]], np.float32)
    out = orc.propagate(ev, np.array([[4, 4, 4], [9, 9, 9]], np.float32), rec, tgt)
    # t0 = 1 + .5*4 = 3 ; t1 = 10 + 2*3 = 16 ; t2 = 100 + .25*16 = 104
    assert np.array_equal(out, [[3] * 3, [16] * 3, [104] * 3])
    ev["radianceMask"] = [0.0, 1.0]                                                # unbiased suffix: queried radiance masked off
    assert np.array_equal(orc.propagate(ev, np.array([[4, 4, 4], [9, 9, 9]], np.float32), rec, tgt),
                          [[1] * 3, [12] * 3, [103] * 3])


def test_accumulate_modes(orc):
    r = np.array([[1, 2, 3]], np.float32); t = np.array([[0.5, 0.5, 2]], np.float32)
    o = np.array([[10, 10, 10, 0]], np.float32)
    assert np.array_equal(orc.accumulate(r, t, o, 0, 3), [[10.125, 10.25, 11.5, 1]])      # += T*R/(iter+1)
    assert np.array_equal(orc.accumulate(r, t, o, 2, 3), [[0.5, 1, 6, 1]])
    assert np.array_equal(orc.accumulate(r, t, o, 4, 3), [[1, 2, 3, 1]])
    assert np.array_equal(orc.accumulate(r, t, o, 5, 3), [[0.5, 0.5, 2, 1]])
    assert np.array_equal(orc.accumulate(r, t, o, 1, 3), o) and np.array_equal(orc.accumulate(r, t, o, 3, 3), o)


def test_adjust_tile_size(orc, pkg):
    for tx, ty, n in [(8, 8, 65536), (8, 8, 30000), (2, 2, 10), (13, 7, 200000), (8, 8, 0)]:
        assert orc.adjust_tile_size(tx, ty, n) == pkg.adjust_tile_size(tx, ty, n)
    assert orc.adjust_tile_size(8, 8, 65536) == (9, 9)        # sqrt(1.25)*8 + .5 = 9.44
    assert orc.adjust_tile_size(8, 8, 0) == (2, 2)            # floor at 2


# ------------------------------------------------------------------ GPU
@pytest.mark.gpu
@pytest.mark.parametrize("mode", [0, 1, 2, 3, 4, 5])
def test_gpu_accumulate_bit_exact(pkg, orc, mode):
    import torch
    rng = np.random.default_rng(mode)
    n = 2560 * 3 + 17
    r = rng.uniform(0, 4, (n, 3)).astype(np.float32); t = rng.uniform(0, 1, (n, 3)).astype(np.float32)
    o = rng.uniform(0, 2, (n, 4)).astype(np.float32)
    od = torch.from_numpy(o).cuda()
    pkg.accumulate_render_radiance(torch.from_numpy(r).cuda(), torch.from_numpy(t).cuda(), od, n, mode, 7)
    torch.cuda.synchronize()
    assert np.array_equal(od.cpu().numpy(), orc.accumulate(r, t, o, mode, 7))


@pytest.mark.gpu
def test_gpu_propagate_bit_exact(pkg, orc):
    import torch
    rng = np.random.default_rng(3)
    n_tiles, n_rec = 57600, 65536          # 2K frame: 320 x 180 tiles
    ev, rec, used = _chains(orc, n_tiles, n_rec, rng)
    assert used > 60000
    end_rad = rng.uniform(0, 3, (n_tiles, 3)).astype(np.float32)
    tgt = rng.uniform(0, 1, (n_rec, 3)).astype(np.float32)
    td = torch.from_numpy(tgt).cuda()
    evd = torch.from_numpy(ev.view(np.int32).reshape(-1, 4)).cuda()
    recd = torch.from_numpy(rec.view(np.int32).reshape(-1, 7)).cuda()
    pkg.propagate_train_radiance(evd, torch.from_numpy(end_rad).cuda(), recd, td, n_tiles)
    torch.cuda.synchronize()
    assert np.array_equal(td.cpu().numpy(), orc.propagate(ev, end_rad, rec, tgt))


@pytest.mark.gpu
@pytest.mark.parametrize("n_rec", [65536, 40000, 777, 1, 100000])
def test_gpu_permute_bit_exact(pkg, orc, synth, n_rec):
    import torch
    q = synth.cornell_queries(65536, seed=2)
    t = synth.radiance_targets(q, seed=2)
    qd, td = torch.zeros(65536, 15, device="cuda"), torch.zeros(65536, 3, device="cuda")
    pkg.permute_train_data(torch.from_numpy(q).cuda(), torch.from_numpy(t).cuda(), qd, td, n_rec, 99)
    torch.cuda.synchronize()
    rq, rt = orc.permute(q, t, n_rec, 99)
    assert np.array_equal(qd.cpu().numpy(), rq) and np.array_equal(td.cpu().numpy(), rt)


@pytest.mark.gpu
@pytest.mark.parametrize("enc", [0, 1])
def test_gpu_fused_shuffle_equals_permute_then_train(pkg, synth, enc):
    """train_shuffled(b) must leave exactly the weights that permute_train_data + train(dst + b*16384) leaves."""
    import torch
    q = synth.cornell_queries(50000, seed=4)
    t = synth.radiance_targets(q, seed=4)
    qs, ts = torch.from_numpy(q).cuda(), torch.from_numpy(t).cuda()
    res = []
    for fused in (False, True):
        net = pkg.Network()
        net.init(0, enc)
        losses = []
        if fused:
            for b in range(4):
                losses.append(net.train_shuffled(qs, ts, 50000, 11, b, want_loss=True))
        else:
            qd, td = torch.zeros(65536, 15, device="cuda"), torch.zeros(65536, 3, device="cuda")
            pkg.permute_train_data(qs, ts, qd, td, 50000, 11)
            for b in range(4):
                losses.append(net.train(qd[b * 16384:], td[b * 16384:], want_loss=True))
        res.append((np.array(losses), net.get_params(pkg.PARAMS_MASTER_F32), net.get_params(pkg.PARAMS_EMA_F16)))
        net.destroy()
    if enc == 0:      # deterministic gradient reduction: bit-identical
        assert np.array_equal(res[0][0], res[1][0]) and np.array_equal(res[0][1], res[1][1]) and np.array_equal(res[0][2], res[1][2])
    else:             # Hash: grid gradients use atomics
        assert np.allclose(res[0][0], res[1][0], rtol=1e-3)


@pytest.mark.gpu
def test_gpu_update_matches_explicit_sequence(pkg, orc, synth):
    """nrc_b200_update == infer -> accumulate -> propagate -> permute -> 4 x train (Device.cpp:2495-2515), two frames."""
    import ctypes
    import torch
    rng = np.random.default_rng(8)
    n_px, n_tiles, n_rec = 64 * 48, (64 // 8) * (48 // 8), 3000
    outs = []
    for use_update in (False, True):
        net = pkg.Network()
        net.init(0, 0)
        out4 = torch.zeros(n_px, 4, device="cuda")
        for frame in range(2):
            frng = np.random.default_rng(100 + frame)
            q = torch.from_numpy(synth.cornell_queries(n_px + n_tiles, seed=frame)).cuda()
            res = torch.zeros(n_px + n_tiles, 3, device="cuda")
            thr = torch.from_numpy(frng.uniform(0, 1, (n_px, 3)).astype(np.float32)).cuda()
            ev, rec, used = _chains(orc, n_tiles, 65536, frng, max_len=120)
            evd = torch.from_numpy(ev.view(np.int32).reshape(-1, 4)).cuda()
            recd = torch.from_numpy(rec.view(np.int32).reshape(-1, 7)).cuda()
            tq_np = synth.cornell_queries(65536, seed=50 + frame)
            tq = torch.from_numpy(tq_np).cuda()
            tt = torch.from_numpy(synth.radiance_targets(tq_np, seed=50 + frame)).cuda()
            if use_update:
                f = pkg.Frame(q.data_ptr(), res.data_ptr(), n_px, n_tiles, thr.data_ptr(), out4.data_ptr(), pkg.RENDER_FULL, frame,
                              evd.data_ptr(), recd.data_ptr(), tq.data_ptr(), tt.data_ptr(), n_rec, frame)
                stat = net.update(f, want_stat=True)
                loss = stat.loss
                assert stat.num_train_records == n_rec
            else:
                net.infer(q, res, n_px + n_tiles)
                pkg.accumulate_render_radiance(res, thr, out4, n_px, pkg.RENDER_FULL, frame)
                pkg.propagate_train_radiance(evd, res[n_px:], recd, tt, n_tiles)
                qd, td = torch.zeros(65536, 15, device="cuda"), torch.zeros(65536, 3, device="cuda")
                pkg.permute_train_data(tq, tt, qd, td, n_rec, frame)
                loss = np.float32(0)
                for b in range(4):
                    loss += np.float32(net.train(qd[b * 16384:], td[b * 16384:], want_loss=True))
                loss *= np.float32(0.25)
            torch.cuda.synchronize()
        outs.append((out4.cpu().numpy(), net.get_params(pkg.PARAMS_MASTER_F32), float(loss), res.cpu().numpy()))
        net.destroy()
    assert np.array_equal(outs[0][0], outs[1][0])          # accumulated image
    assert np.array_equal(outs[0][1], outs[1][1])          # weights after 8 optimizer steps
    assert np.array_equal(outs[0][3], outs[1][3])
    assert abs(outs[0][2] - outs[1][2]) <= 1e-6 * abs(outs[0][2])
    assert (outs[0][0][:, :3] != 0).any()                  # frame 2 sees a trained cache (EMA no longer zero)


# ------------------------------------------------------------------ pinned by the REFERENCE's own helper kernels
# tests/golden/ref_helpers.npz, ref_frame_*.npz: outputs of the unmodified nrc/shaders/nrc_helpers.cu kernels (and, for the
# frame, the reference network) run on a B200 by tools/gen_golden_ref.py; inputs are regenerated from seeds (golden_util).
import golden_util as gu  # noqa: E402


def _helpers():
    g = gu.load("ref_helpers.npz")
    if g is None:
        pytest.skip("tests/golden/ref_helpers.npz missing (generate with tools/gen_golden_ref.py on a B200)")
    return g, gu.helper_inputs()


def _structured(orc, ev, rec):
    return (np.ascontiguousarray(ev).view(orc.END_VERTEX_DTYPE).reshape(-1), np.ascontiguousarray(rec).view(orc.RECORD_DTYPE).reshape(-1))


def test_oracle_helpers_match_reference_kernels(orc):
    """The oracle's restatement of accumulate / propagate / permute against the reference kernels' outputs, bit for bit."""
    g, h = _helpers()
    for mode, it in gu.ACC_CASES:
        assert np.array_equal(orc.accumulate(h["acc_radiance"], h["acc_throughput"], h["acc_output"], mode, it), g[f"acc_m{mode}_i{it}"]), (mode, it)
    ev, rec = _structured(orc, h["prop_ev"], h["prop_rec"])
    assert np.array_equal(orc.propagate(ev, h["prop_end_radiance"], rec, h["prop_targets"]), g["prop_targets"])
    for n_rec in gu.PERMUTE_CASES:      # the gather itself, with the reference-side permutation supplied
        src = h["perm"] % min(n_rec, 65536)
        assert np.array_equal(gu.digest(h["perm_q"][src]), g[f"perm_{n_rec}_q_digest"])
        assert np.array_equal(gu.digest(h["perm_t"][src]), g[f"perm_{n_rec}_t_digest"])
        assert np.array_equal(h["perm_q"][src][::97], g[f"perm_{n_rec}_q_sample"])


@pytest.mark.gpu
def test_gpu_helpers_match_reference_kernels(pkg):
    """The CUDA helper kernels, through the C ABI, against the reference kernels' outputs: bit-exact."""
    import torch
    g, h = _helpers()
    dev = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
    n_px = gu.HELPER_W * gu.HELPER_H
    for mode, it in gu.ACC_CASES:
        o = dev(h["acc_output"])
        pkg.accumulate_render_radiance(dev(h["acc_radiance"]), dev(h["acc_throughput"]), o, n_px, mode, it)
        torch.cuda.synchronize()
        assert np.array_equal(o.cpu().numpy(), g[f"acc_m{mode}_i{it}"]), (mode, it)
    t = dev(h["prop_targets"])
    pkg.propagate_train_radiance(dev(h["prop_ev"]), dev(h["prop_end_radiance"]), dev(h["prop_rec"]), t, gu.HELPER_TILES_X * gu.HELPER_TILES_Y)
    torch.cuda.synchronize()
    assert np.array_equal(t.cpu().numpy(), g["prop_targets"])
    for n_rec in gu.PERMUTE_CASES:
        qd, td = torch.zeros(65536, 15, device="cuda"), torch.zeros(65536, 3, device="cuda")
        pkg.permute_train_data_with(dev(h["perm_q"]), dev(h["perm_t"]), qd, td, dev(h["perm"]), n_rec)
        torch.cuda.synchronize()
        assert np.array_equal(gu.digest(qd.cpu().numpy()), g[f"perm_{n_rec}_q_digest"])
        assert np.array_equal(gu.digest(td.cpu().numpy()), g[f"perm_{n_rec}_t_digest"])


@pytest.mark.gpu
@pytest.mark.parametrize("enc,name", [(0, "freq"), (1, "hash")])
def test_gpu_update_matches_reference_frame(pkg, enc, name):
    """nrc_b200_update against the REFERENCE's frame sequence (its network + its helper kernels, Device.cpp:2495-2515) on
    the same seeded frame: radiance / accumulated image / propagated targets at the radiance tolerance, the shuffled batches
    byte for byte (digest), the four training losses at the loss-curve tolerance (2 %)."""
    import torch
    g = gu.load(f"ref_frame_{name}.npz")
    if g is None:
        pytest.skip("tests/golden/ref_frame_*.npz missing (generate with tools/gen_golden_ref.py on a B200)")
    f = gu.frame_inputs(enc)
    dev = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
    rel = lambda a, b, floor: float((np.abs(a - b) / np.maximum(np.abs(b), floor)).max())
    res = {}
    for mode in (pkg.ACC_FP32, pkg.ACC_FP16):
        net = pkg.Network()
        net.init(0, enc)
        net.load_master(gu.seeded_master(enc, net.n_params()[0], True))
        q0, t0 = gu.batch(0)
        lr = net.getLearningRate()
        net.setHyperParams(pkg.HyperParams(0.0))         # the fixture's warm-up: weights -> EMA, Adam moments move
        net.train(dev(q0), dev(t0))
        net.setHyperParams(pkg.HyperParams(lr))
        net.set_accumulator(mode)
        n_px, (tx, ty) = f["n_px"], f["tiles"]
        fq, out3 = dev(f["queries"]), torch.zeros((f["queries"].shape[0], 3), device="cuda")
        thr, out4 = dev(f["throughput"]), dev(f["output"])
        ev, rec, tq, tt = dev(f["ev"]), dev(f["rec"]), dev(f["train_q"]), dev(f["train_t"])
        fr = pkg.Frame(fq.data_ptr(), out3.data_ptr(), n_px, tx * ty, thr.data_ptr(), out4.data_ptr(), pkg.RENDER_FULL, 2,
                       ev.data_ptr(), rec.data_ptr(), tq.data_ptr(), tt.data_ptr(), f["used"], gu.FRAME_SEED)
        stat = net.update(fr, want_stat=True)
        torch.cuda.synchronize()
        r = {"results": rel(out3.cpu().numpy()[::7], g["results"], 0.05),
             "output": rel(out4.cpu().numpy()[::5][:, :3], g["output"][:, :3], 0.05),
             "targets": rel(tt.cpu().numpy()[:f["used"]][::3], g["targets"], 0.05),
             "loss": abs(stat.loss - float(g["losses"].mean())) / float(g["losses"].mean())}
        # the shuffled batches the training steps consumed: materialise them with the stand-alone permute and compare bytes
        qd, td = torch.zeros(65536, 15, device="cuda"), torch.zeros(65536, 3, device="cuda")
        pkg.permute_train_data(tq, dev(f["train_t"]), qd, td, f["used"], gu.FRAME_SEED)
        torch.cuda.synchronize()
        assert np.array_equal(gu.digest(qd.cpu().numpy()), g["permuted_q_digest"])
        res[mode] = r
        print(f"[frame {name}, acc mode {mode}] max rel (floor 0.05): radiance {r['results']:.2e}, image {r['output']:.2e}, "
              f"propagated targets {r['targets']:.2e}; mean loss of the 4 steps off by {100 * r['loss']:.3f} %")
        assert (out4.cpu().numpy()[:, 3] == 1.0).all()
        assert r["results"] <= 1e-2 and r["output"] <= 1e-2 and r["targets"] <= 1e-2, r
        assert r["loss"] <= 0.02, r
        net.destroy()
