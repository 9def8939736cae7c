"""nrc_b200_train_grad_noalloc (include/nrc_b200.h): the gradient-only training step with its scratch in the library's static
__device__ array -- no counterpart in the reference.  Checked through the C ABI
against the existing entry point nrc_b200_train_grad on the same weights and inputs, on two calls in a row (the second one
with a smaller batch, so stale scratch contents of the first call lie behind it).  Collected last on purpose (zz)."""
import ctypes
import importlib

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _read_device_float(ptr):
    out = np.empty(1, dtype=np.float32)
    rt = ctypes.CDLL("libcudart.so.12")
    rt.cudaMemcpy.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_size_t, ctypes.c_int]
    assert rt.cudaMemcpy(out.ctypes.data, ptr, 4, 2) == 0      # cudaMemcpyDeviceToHost (synchronises)
    return float(out[0])


def _reference_grad(pkg, torch, enc, q_d, t_d, n):
    """Existing entry point on a FRESH handle with the same seed (= same weights): train_grad leaves a pending gradient and,
    Hash, a dirty scatter accumulator behind, so every comparison gets its own handle."""
    ref = pkg.Network()
    ref.init(0, enc)
    ref.train_grad(q_d, t_d, n, n)
    torch.cuda.synchronize()
    g = ref.get_params(pkg.GRAD_F32).copy()
    loss_p = ctypes.c_void_p()
    assert pkg.lib().nrc_b200_last_loss_device(ref._h, ctypes.byref(loss_p)) == 0
    loss = _read_device_float(loss_p.value)
    ref.destroy()
    return g, loss


@pytest.mark.parametrize("enc", [0, 1], ids=["frequency", "hash"])
def test_train_grad_noalloc_matches_train_grad_on_two_calls_in_a_row(pkg, enc):
    import torch
    synth = importlib.import_module("nrc_b200.synthetic")
    max_samples, scratch_bytes = pkg.noalloc_limits()
    assert max_samples == 4096 and 0 < scratch_bytes <= 8 << 20
    tq = synth.cornell_queries(max_samples, seed=21)
    tt = synth.radiance_targets(tq, seed=21)
    q_d, t_d = torch.from_numpy(tq).cuda(), torch.from_numpy(tt).cuda()
    net = pkg.Network()
    net.init(0, enc)
    n_params, n_mlp = net.n_params()
    grad_d = torch.full((n_params,), float("nan"), device="cuda")
    loss_d = torch.full((1,), float("nan"), device="cuda")
    before = net.get_params(pkg.PARAMS_MASTER_F32).copy()
    for n, off in ((max_samples, 0), (1024, 512)):          # two calls in a row on the same handle, sharing the static scratch
        qs, ts = q_d[off:off + n].contiguous(), t_d[off:off + n].contiguous()
        grad_d.fill_(float("nan"))
        torch.cuda.synchronize()
        mem_before = torch.cuda.mem_get_info()[0]
        net.train_grad_noalloc(qs, ts, n, n, grad_d, loss_d)
        torch.cuda.synchronize()
        if off:   # second call: kernels and the scratch symbol are loaded (CUDA loads a module lazily at its first launch)
            assert torch.cuda.mem_get_info()[0] == mem_before, "the call changed the amount of free device memory"
        g, loss = grad_d.cpu().numpy(), float(loss_d.item())
        g_ref, loss_ref = _reference_grad(pkg, torch, enc, qs, ts, n)
        assert np.isfinite(g).all() and np.abs(g_ref).max() > 0
        # weight gradients: fixed-order fp32 sums of per-tile tensor-core partials -> bit-identical to the existing entry point
        assert np.array_equal(g[:n_mlp], g_ref[:n_mlp]), f"MLP gradient differs (n={n})"
        assert abs(loss - loss_ref) <= 1e-6 * abs(loss_ref), f"loss {loss} vs {loss_ref}"
        if enc == 1:   # grid gradients are fp32 atomics (order-dependent): the tolerance of tests/test_train_gpu.py
            gg, gg_ref = g[n_mlp:], g_ref[n_mlp:]
            assert np.abs(gg - gg_ref).max() <= 1e-2 * np.abs(gg_ref).max(), f"grid gradient differs (n={n})"
            # same set of touched entries (a sum of atomics may cancel to exactly 0 in one order and not in the other)
            assert abs(np.count_nonzero(gg) - np.count_nonzero(gg_ref)) <= 1e-3 * np.count_nonzero(gg_ref)
    assert np.array_equal(net.get_params(pkg.PARAMS_MASTER_F32), before), "a gradient-only call changed the weights"
    # the handle is still good for an ordinary optimizer step afterwards (grid-barrier bookkeeping stayed consistent)
    tq2 = synth.cornell_queries(pkg.BATCH_SIZE, seed=22)
    tt2 = synth.radiance_targets(tq2, seed=22)
    loss2 = net.train(torch.from_numpy(tq2).cuda(), torch.from_numpy(tt2).cuda(), want_loss=True)
    assert np.isfinite(loss2) and loss2 > 0
    net.destroy()


def test_train_grad_noalloc_rejects_sizes_beyond_the_static_scratch(pkg):
    import torch
    max_samples, _ = pkg.noalloc_limits()
    net = pkg.Network()
    net.init(0, pkg.FREQUENCY)
    n = max_samples + 128
    q_d, t_d = torch.zeros(n, 15, device="cuda"), torch.zeros(n, 3, device="cuda")
    grad_d = torch.zeros(net.n_params()[0], device="cuda")
    with pytest.raises(ValueError, match="exceeds the static scratch maximum of 4096 samples"):
        net.train_grad_noalloc(q_d, t_d, n, n, grad_d)
    with pytest.raises(ValueError, match="multiple of 128"):
        net.train_grad_noalloc(q_d, t_d, 100, 100, grad_d)
    net.destroy()
