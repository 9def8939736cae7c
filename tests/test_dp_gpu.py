"""2-GPU test of the data-parallel training paths (run with `gpurun --gpus 2`; skipped with fewer devices):
both ranks must end with bit-identical weights, the fused in-kernel all-reduce must match the NCCL path, and the
global loss must match a single-GPU step on the concatenated batch."""
import os
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, out, mode, enc):
    sys.path.insert(0, ROOT)
    import importlib
    import torch
    import torch.distributed as dist
    import __graft_entry__ as entry
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world,
                            device_id=torch.device("cuda", rank))
    pkg = entry.load_package()
    synth = entry.load_package_synthetic()
    dp = importlib.import_module("nrc_b200.dist")
    net = pkg.Network()
    net.init(torch.cuda.current_stream().cuda_stream, enc)
    trainer = dp.DataParallelTrainer(net, mode=mode)
    n_local = 8192
    losses = []
    for step in range(6):
        q = synth.cornell_queries(n_local * world, seed=100 + step)
        t = synth.radiance_targets(q, seed=100 + step)
        lo, hi = rank * n_local, (rank + 1) * n_local
        qd, td = torch.from_numpy(q[lo:hi]).cuda(), torch.from_numpy(t[lo:hi]).cuda()
        losses.append(trainer.train(qd, td, n_local, torch.cuda.current_stream().cuda_stream, want_loss=True))
    np.savez(f"{out}.{mode}.{rank}.npz", master=net.get_params(pkg.PARAMS_MASTER_F32), ema=net.get_params(pkg.PARAMS_EMA_F16),
             losses=np.array(losses))
    net.destroy()
    dist.destroy_process_group()


@pytest.mark.parametrize("enc", [0, 1])
def test_dp_fused_matches_nccl_and_single_gpu(pkg, synth, tmp_path, enc):
    import torch
    import torch.multiprocessing as mp
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    world = 2
    res = {}
    for k, mode in enumerate(("nccl", "fused")):
        out = str(tmp_path / f"dp{enc}")
        mp.spawn(_worker, args=(world, 29600 + 10 * enc + k + os.getpid() % 200, out, mode, enc), nprocs=world, join=True)
        res[mode] = [np.load(f"{out}.{mode}.{r}.npz") for r in range(world)]
        a, b = res[mode]
        assert np.array_equal(a["master"], b["master"]) and np.array_equal(a["ema"], b["ema"]), f"{mode}: replicas differ"
        assert np.allclose(a["losses"], b["losses"], rtol=1e-6)
    # single GPU on the concatenated batch (same global batch of 16384)
    net = pkg.Network()
    net.init(0, enc)
    ref = []
    for step in range(6):
        q = synth.cornell_queries(16384, seed=100 + step)
        t = synth.radiance_targets(q, seed=100 + step)
        ref.append(net.train(torch.from_numpy(q).cuda(), torch.from_numpy(t).cuda(), want_loss=True))
    single = net.get_params(pkg.PARAMS_MASTER_F32)
    net.destroy()
    lr = 1e-3 if enc == 0 else 1e-2
    nm = 22528 if enc == 0 else 21504
    for mode in ("nccl", "fused"):
        assert np.allclose(res[mode][0]["losses"], ref, rtol=2e-3), (mode, res[mode][0]["losses"], ref)
        # 6 steps move a weight by at most 6 lr.  Data parallel and single GPU sum the same 128 per-tile partials in a different
        # association (rank by rank vs strided groups), so they agree to a small fraction of ONE step: measured <= 0.03 lr
        # (Frequency) / 0.1 lr (Hash: fp32 atomics in the grid gradient) at the maximum, bit-identical for most weights.
        d = np.abs(res[mode][0]["master"][:nm] - single[:nm])
        print(f"[dp {mode} enc {enc}] max |w_dp - w_single| = {d.max():.3e} ({d.max() / lr:.4f} lr), identical {np.mean(d == 0):.3f}, "
              f"within 0.01 lr {np.mean(d <= 0.01 * lr):.4f}")
        assert d.max() <= 0.25 * lr and np.mean(d <= 0.02 * lr) > 0.99
    if enc == 0:   # same per-rank partial sums, added in rank order by both exchanges: bit-identical
        assert np.array_equal(res["nccl"][0]["master"], res["fused"][0]["master"])
        assert np.array_equal(res["nccl"][0]["ema"], res["fused"][0]["ema"])
