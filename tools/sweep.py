"""BASELINE.json configs[4]: throughput sweep of query batch sizes 2^14..2^24 (both encodings) and HashGrid table sizes
2^14..2^22 entries per level, one GPU (or every rank of a torchrun job: shards are independent, no collective),
against the rooflines of DESIGN.md section 3.

    python tools/sweep.py [--out profiles/r01_sweep.json] [--max-log2-batch 24]
    python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 tools/sweep.py   (weak: every rank a full batch)

Per point: CUDA-event time of `reps` back-to-back launches after warm-up, inputs resident in HBM.  Batches whose inputs fit in
the 126 MB L2 (< 2^21 queries) are timed with an L2 flush (a 256 MB memset) between launches, the flush excluded by timing each
launch with its own event pair.  Reported per point: us per launch, G queries/s, TFLOP/s of the padded MLP (2 * P_mlp per
query) and its fraction of MEASURED_PEAKS.json's tensor peak, algorithmic HBM GB/s (72 B/query), and for Hash the gather
bytes through L1/L2 (512 B/query).
"""
import argparse
import importlib
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as entry  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "sweep.json"))
    ap.add_argument("--min-log2-batch", type=int, default=14)
    ap.add_argument("--max-log2-batch", type=int, default=24)
    ap.add_argument("--reps", type=int, default=20)
    args = ap.parse_args()
    import torch
    import torch.distributed as dist
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("tools/sweep.py needs a B200 (there is no CPU path)")
    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    pkg = entry.load_package()
    synth = importlib.import_module("nrc_b200.synthetic")
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak_tf, peak_hbm = peaks.get("bf16_tflops", 1590.0), peaks.get("hbm_gbs", peaks.get("hbm_gb_s", 6500.0))
    stream = torch.cuda.current_stream()
    sp = stream.cuda_stream
    n_max = 1 << args.max_log2_batch
    base = synth.cornell_queries(1 << 16, seed=11 + rank)           # incoherent surface points (worst case for the grid)
    q_d = torch.from_numpy(base).cuda().repeat((n_max >> 16) or 1, 1)[:n_max].contiguous()
    # decorrelate the tiled copies a little so that the hash-grid gathers do not repeat every 65536 queries
    q_d[:, :3] = (q_d[:, :3] + 0.05 * (torch.rand((n_max, 3), device="cuda") - 0.5)).clamp_(-0.05, 0.05)
    out_d = torch.empty((n_max, 3), device="cuda")
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")

    def time_infer(net, n, reps):
        for _ in range(3):
            net.infer(q_d, out_d, n, sp)
        torch.cuda.synchronize()
        small = n * 72 < (126 << 20)
        if small:                                                    # flush L2 between launches, time each launch alone
            ts = []
            for _ in range(reps):
                flush.zero_()
                a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a.record(stream)
                net.infer(q_d, out_d, n, sp)
                b.record(stream)
                torch.cuda.synchronize()
                ts.append(a.elapsed_time(b))
            return float(np.median(ts)), "L2 flushed between launches, median of single launches"
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(stream)
        for _ in range(reps):
            net.infer(q_d, out_d, n, sp)
        b.record(stream)
        torch.cuda.synchronize()
        return a.elapsed_time(b) / reps, "inputs larger than L2, back-to-back launches"

    def point(enc, log2_table, n, ms, how):
        e_w = 80 if enc == 0 else 64
        flop = 2 * (64 * e_w + 4 * 64 * 64 + 16 * 64)
        t = n * world
        if world > 1:
            tt = torch.tensor([ms], device="cuda")
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            ms_ = float(tt.item())
        else:
            ms_ = ms
        d = {"encoding": "frequency" if enc == 0 else "hash", "log2_table": log2_table, "n_gpus": world,
             "queries_per_gpu": n, "us_per_launch": ms_ * 1e3, "gq_per_s": t / ms_ / 1e6,
             "tflops": t * flop / ms_ / 1e9, "frac_tensor_peak": t * flop / ms_ / 1e9 / (peak_tf * world),
             "hbm_gbs_algorithmic": t * 72 / ms_ / 1e6, "frac_hbm_peak": t * 72 / ms_ / 1e6 / (peak_hbm * world), "timing": how}
        if enc == 1:
            d["gather_gbs"] = t * 512 / ms_ / 1e6
        return d

    rows = []
    # (1) batch-size sweep, both encodings (Hash at the reference's 2^15 table)
    for enc in (0, 1):
        net = pkg.Network()
        net.init(sp, enc)
        if enc == 1:     # trained-like grid values instead of the +-1e-4 initialisation (same traffic, non-trivial numbers)
            master = net.get_params(pkg.PARAMS_MASTER_F32)
            net.set_params(pkg.PARAMS_EMA_F16, master.astype(np.float16))
        for lg in range(args.min_log2_batch, args.max_log2_batch + 1):
            n = 1 << lg
            ms, how = time_infer(net, n, args.reps)
            rows.append(point(enc, 15 if enc == 1 else None, n, ms, how))
        net.destroy()
    # (2) table-size sweep (Hash), at the 2K frame's query count and at 2^20
    for lt in range(14, 23):
        net = pkg.Network()
        net.init(sp, 1, log2_hashmap_size=lt)
        for n in (1 << 20, 3744000):
            if n > n_max:
                continue
            ms, how = time_infer(net, n, max(5, args.reps // 2))
            r = point(1, lt, n, ms, how)
            r["table_bytes"] = int(net.n_params()[0] - net.n_params()[1]) * 2
            rows.append(r)
        # one training step at this table size (optimizer pass over the whole table: 26 B/param)
        tq = synth.cornell_queries(pkg.BATCH_SIZE, seed=4)
        tt = synth.radiance_targets(tq, seed=4)
        tq_d, tt_d = torch.from_numpy(tq).cuda(), torch.from_numpy(tt).cuda()
        for _ in range(3):
            net.train(tq_d, tt_d, sp)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        a.record(stream)
        for _ in range(10):
            net.train(tq_d, tt_d, sp)
        b.record(stream)
        torch.cuda.synchronize()
        us = a.elapsed_time(b) / 10 * 1e3
        n_grid = int(net.n_params()[0] - net.n_params()[1])
        rows.append({"encoding": "hash", "log2_table": lt, "n_gpus": world, "train_step_us": us, "train_samples_per_s": pkg.BATCH_SIZE / us * 1e6,
                     "optimizer_state_gbs": n_grid * 26 / us / 1e3, "grid_params": n_grid})
        net.destroy()
    if rank == 0:
        os.makedirs(os.path.dirname(args.out), exist_ok=True)
        json.dump({"peaks": {"tensor_tflops": peak_tf, "hbm_gbs": peak_hbm}, "rows": rows}, open(args.out, "w"), indent=1)
        for r in rows:
            if "us_per_launch" in r:
                print(f"{r['encoding']:9s} table 2^{r['log2_table']}  n={r['queries_per_gpu']:>9d}  {r['us_per_launch']:9.1f} us  {r['gq_per_s']:7.2f} Gq/s  "
                      f"{r['tflops']:7.1f} TF/s ({r['frac_tensor_peak']:.2f})  {r['hbm_gbs_algorithmic']:7.0f} GB/s ({r['frac_hbm_peak']:.2f})")
            else:
                print(f"hash      table 2^{r['log2_table']}  train step {r['train_step_us']:8.1f} us  optimizer state {r['optimizer_state_gbs']:7.0f} GB/s")
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
