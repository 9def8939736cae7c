#!/usr/bin/env python
"""bench.py -- the NRC update benchmark (BASELINE.json metric) on N B200s of one node.

A "step" is ONE NRC UPDATE of a 2K frame, the per-frame hot path of the reference's Device::render
(/root/reference/nrc/src/Device.cpp:2466-2515):
    Network::infer on (#px + #tiles) = 2560*1440 + 320*180 = 3 744 000 RadianceQuery records, then
    NUM_BATCHES = 4 sequential Network::train steps of BATCH_SIZE = 16384 path-tail samples each.
Frequency encoding (the reference's default, Device.cpp:274) unless --encoding hash.

value      = (queries + training samples) processed per second over all ranks, inputs resident in HBM.
e2e        = the same through the host-buffer C-ABI calls (pinned host memory in, results back to host).
roofline   = the dominant kernel (fused query kernel) against the measured tensor peak.
cpu_baseline / --impl reference = the scalar CPU oracle (the reference has no CPU path; oracle/ is the port).

Multi-GPU (torchrun): every rank owns a full 2K frame of queries (batch-sharded inference, no collective) and
16384 local training samples per step; the fp32 weight gradients are summed with an NCCL all-reduce between the
backward kernel and the Adam+EMA kernel (global batch = N*16384), so per-GPU work is fixed: "scaling": "weak".
"""
import argparse
import importlib
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
import __graft_entry__ as entry  # noqa: E402

METRIC = "cache queries/s + train samples/s (5x64 MLP); ms per 2K-frame NRC update"
UNIT = "samples/s"               # queries + training samples per second
FRAME_QUERIES = 2560 * 1440 + (2560 // 8) * (1440 // 8)      # 3 744 000 (Device.cpp:1275: 8x8 tiles)
NUM_BATCHES, BATCH_SIZE = 4, 16384
FLOP_PER_QUERY = {0: 45056, 1: 43008}                          # 2*P_mlp
FLOP_PER_TRAIN_SAMPLE = {0: 124928, 1: 129024}


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", choices=["ours", "reference"], default="ours")
    ap.add_argument("--encoding", choices=["frequency", "hash"], default="frequency")
    ap.add_argument("--queries", type=int, default=FRAME_QUERIES, help="queries per update per GPU")
    ap.add_argument("--query-order", choices=["frame", "random"], default="frame",
                    help="frame: pixel-ordered primary hits of a camera in the Cornell box + incoherent tile tails (what a "
                         "renderer hands over); random: independent random surface points (incoherent worst case)")
    ap.add_argument("--per-batch-launch", action="store_true",
                    help="one nrc_b200_train launch per batch (the reference's call pattern) instead of nrc_b200_train_batches")
    ap.add_argument("--no-other-encoding", action="store_true", help="skip the short run of the other encoding")
    ap.add_argument("--shard-frame", type=int, default=0, metavar="TOTAL",
                    help="strong scaling (BASELINE.json configs[3]): ONE frame of TOTAL queries (4K: 8424000) batch-sharded over the "
                         "ranks, 16384-sample training batches split across the ranks; default: every rank owns a full frame")
    ap.add_argument("--e2e-steps", type=int, default=10)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--dp", choices=["fused", "nccl"], default="fused",
                    help="multi-GPU gradient all-reduce: fused into the step kernel over NVLink peer memory, or NCCL")
    ap.add_argument("--ref-gpu", action="store_true",
                    help="also time the reference's own tcnn build (oracle/_ref/libnrc_ref.so) on this GPU")
    return ap.parse_args()


def workload_name(enc, nq):
    return (f"2K-frame NRC update: infer {nq} queries (2560x1440 px + 320x180 tiles) + {NUM_BATCHES}x{BATCH_SIZE} "
            f"training samples, {'Frequency' if enc == 0 else 'Hash'} encoding, 5x64 ReLU MLP")


def make_inputs(synth, nq, rank, order="frame"):
    """Seeded Cornell-box queries (host, float32).  order == "frame": 2560-pixel-wide rows of primary hits in pixel order,
    the last 320x180 (or 1/65 of a non-default count) being incoherent tile tails; "random": independent surface points,
    tiled from 2^20-query blocks to keep start-up short."""
    if order == "frame":
        n_tail = 320 * 180 if nq == FRAME_QUERIES else nq // 65
        rows = -(-(nq - n_tail) // 2560)
        px = synth.cornell_frame(2560, rows, seed=1337 + 97 * rank)[:nq - n_tail]
        q = np.concatenate([px, synth.cornell_queries(n_tail, seed=1342 + 97 * rank)], axis=0) if n_tail else px
    else:
        blocks, left, b = [], nq, 0
        while left > 0:
            m = min(left, 1 << 20)
            blocks.append(synth.cornell_queries(m, seed=1337 + 97 * rank + b))
            left -= m
            b += 1
        q = np.concatenate(blocks, axis=0)
    tq = synth.cornell_queries(NUM_BATCHES * BATCH_SIZE, seed=4242 + rank)
    tt = synth.radiance_targets(tq, seed=4242 + rank)
    return np.ascontiguousarray(q), tq, tt


class ClockSampler(threading.Thread):
    """Samples SM clock + throttle reasons of one GPU through NVML while the timed region runs."""
    BITS = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown", 0x4: "sw_power_cap",
            0x80: "hw_power_brake"}

    def __init__(self, index, period=0.002):
        super().__init__(daemon=True)
        self.index, self.period = index, period
        self.samples, self.reasons, self.max_mhz, self.err = [], set(), None, None
        self._stop_evt = threading.Event()
        self._nvml, self._h = None, None
        try:   # NVML is brought up HERE (main thread, before the timed region): nvmlInit alone outlasts a 60 ms region
            import pynvml
            pynvml.nvmlInit()
            # NVML enumerates physical devices; honour CUDA_VISIBLE_DEVICES when it is a plain index list
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            idx = self.index
            if vis:
                try:
                    idx = int(vis.split(",")[self.index])
                except ValueError:
                    pass
            self._h = pynvml.nvmlDeviceGetHandleByIndex(idx)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self._h, pynvml.NVML_CLOCK_SM)
            self._nvml = pynvml
        except Exception as e:  # NVML missing: report it, never fake a clock
            self.err = repr(e)

    def sample_now(self):
        """One sample from the calling thread (the bench calls it while the timed launches are still executing)."""
        if self._nvml is None:
            return
        try:
            self.samples.append(self._nvml.nvmlDeviceGetClockInfo(self._h, self._nvml.NVML_CLOCK_SM))
            r = self._nvml.nvmlDeviceGetCurrentClocksThrottleReasons(self._h)
            for bit, name in self.BITS.items():
                if r & bit:
                    self.reasons.add(name)
        except Exception as e:
            self.err = repr(e)

    def run(self):
        while self._nvml is not None and not self._stop_evt.is_set():
            self.sample_now()
            time.sleep(self.period)

    def stop(self):
        self._stop_evt.set()
        self.join(timeout=2)
        d = {"sm_mhz": float(np.median(self.samples)) if self.samples else None, "sm_max_mhz": self.max_mhz,
             "reasons": sorted(self.reasons), "samples": len(self.samples)}
        if self.err:
            d["error"] = self.err
        return d


def cpu_update_rate(orc, synth, enc, nq, threads, seconds_budget=20.0):
    """Times the CPU oracle on a bounded sample of the update and extrapolates to the full update:
    t_update = nq * t_query + 65536 * t_train_sample.  Returns (samples/s, description)."""
    tr = orc.Trainer(enc)
    tr.ema[:] = tr.params                      # query with live weights (EMA starts at zero)
    n_inf = 1 << 16
    q = synth.cornell_queries(n_inf, seed=7)
    t0 = time.perf_counter()
    orc.infer(tr.spec, tr.ema, q[:4096], nthreads=threads)         # calibrate
    t_probe = (time.perf_counter() - t0) / 4096
    n_inf = int(min(n_inf, max(4096, seconds_budget * 0.6 / max(t_probe, 1e-9))))
    t0 = time.perf_counter()
    orc.infer(tr.spec, tr.ema, q[:n_inf], nthreads=threads)
    t_q = (time.perf_counter() - t0) / n_inf
    tq = synth.cornell_queries(BATCH_SIZE, seed=8)
    tt = synth.radiance_targets(tq, seed=8)
    t0 = time.perf_counter()
    tr.train_step(tq, tt, nthreads=threads)
    t_s = (time.perf_counter() - t0) / BATCH_SIZE
    tr.close()
    n_train = NUM_BATCHES * BATCH_SIZE
    t_update = nq * t_q + n_train * t_s
    sample = (f"{n_inf} queries + one {BATCH_SIZE}-sample training step on {threads} threads, extrapolated to "
              f"{nq} queries + {n_train} samples ({t_q * 1e6:.2f} us/query, {t_s * 1e6:.2f} us/train sample)")
    return (nq + n_train) / t_update, t_update, sample


def run_reference(args, enc, out):
    """--impl reference: the reference has no CPU implementation of this path (every function is __global__);
    the timed code is oracle/'s scalar C restatement (kind "port") on all host threads.  Each step is a bounded
    sample of the update."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    orc = entry.load_oracle()
    orc.build()
    synth = entry.load_package_synthetic()
    threads = orc.ncpu()
    tr = orc.Trainer(enc)
    tr.ema[:] = tr.params
    # one step = a bounded sample of the update, sized so that warmup+steps stay within ~2 minutes
    probe = synth.cornell_queries(2048, seed=6)
    t0 = time.perf_counter()
    orc.infer(tr.spec, tr.ema, probe, nthreads=threads)
    t_probe = (time.perf_counter() - t0) / 2048
    n_inf = int(np.clip(100.0 / ((args.steps + args.warmup) * 1.6 * t_probe), 1024, 65536)) // 128 * 128
    n_tr = max(128, n_inf // 8)
    q = synth.cornell_queries(n_inf, seed=7)
    tq = synth.cornell_queries(n_tr, seed=8)
    tt = synth.radiance_targets(tq, seed=8)
    t_q, t_s = [], []
    for it in range(args.warmup + args.steps):
        t0 = time.perf_counter()
        orc.infer(tr.spec, tr.ema, q, nthreads=threads)
        t1 = time.perf_counter()
        tr.train_step(tq, tt, nthreads=threads)
        t2 = time.perf_counter()
        if it >= args.warmup:
            t_q.append((t1 - t0) / n_inf)
            t_s.append((t2 - t1) / n_tr)
    tr.close()
    n_train = NUM_BATCHES * BATCH_SIZE
    t_update = args.queries * float(np.mean(t_q)) + n_train * float(np.mean(t_s))
    v = (args.queries + n_train) / t_update
    sample = (f"per step {n_inf} queries + {n_tr} training samples on {threads} host threads; per-unit times "
              f"extrapolated to the full update ({args.queries} queries + {n_train} samples)")
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": t_update * 1e3, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f16", "data": "synthetic",
            "config": {"workload": workload_name(enc, args.queries)},
            "cpu_baseline": {"value": v, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample},
            "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), file=out, flush=True)


class DevPtr:
    """Zero-copy torch view of a raw device pointer (the C ABI hands out plain pointers)."""

    def __init__(self, ptr, n, typestr="<f4"):
        self.__cuda_array_interface__ = {"shape": (n,), "typestr": typestr, "data": (ptr, False), "version": 2}


def run_ours(args, enc):
    import torch
    import torch.distributed as dist
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a B200: there is no CPU path (use --impl reference for the CPU oracle)")
    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    pkg = entry.load_package()
    synth = entry.load_package_synthetic()
    nq = args.queries
    n_train = NUM_BATCHES * BATCH_SIZE
    strong = args.shard_frame > 0
    train_local = BATCH_SIZE                     # samples of each training batch this rank processes
    if strong:                                   # one frame for the whole job: contiguous tile-aligned shard per rank
        lo, hi = importlib.import_module("nrc_b200.dist").shard_range(args.shard_frame, rank, world) if world > 1 else (0, args.shard_frame)
        nq = hi - lo
        train_local = BATCH_SIZE // world

    # ---- inputs: host (pinned) + device-resident copies
    q_np, tq_np, tt_np = make_inputs(synth, nq, rank, args.query_order)
    q_h = torch.from_numpy(q_np).pin_memory()
    tq_h = torch.from_numpy(tq_np).pin_memory()
    tt_h = torch.from_numpy(tt_np).pin_memory()
    out_h = torch.empty((nq, 3), dtype=torch.float32).pin_memory()
    q_d, tq_d, tt_d = q_h.cuda(), tq_h.cuda(), tt_h.cuda()
    out_d = torch.empty((nq, 3), device="cuda")

    stream = torch.cuda.current_stream()
    sp = stream.cuda_stream
    net = pkg.Network()
    net.init(sp, enc)

    dp = importlib.import_module("nrc_b200.dist").DataParallelTrainer(net, mode=args.dp) if world > 1 else None

    def train_batch(b, want_loss=False):
        qs = q_ptr_t + b * BATCH_SIZE * 60
        ts = t_ptr_t + b * BATCH_SIZE * 12
        if world == 1:
            return net.train(qs, ts, sp, want_loss=want_loss)
        # shard -> gradient (normalised by the global batch) -> all-reduce over NVLink -> Adam+EMA
        # (strong scaling: this rank's slice of the SAME 16384-sample batch)
        off = rank * train_local if strong else 0
        return dp.train(qs + off * 60, ts + off * 12, train_local, sp, want_loss=want_loss)

    q_ptr_t, t_ptr_t = tq_d.data_ptr(), tt_d.data_ptr()
    ev_inf = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]

    def update(step_idx=None):
        if step_idx is not None:
            ev_inf[step_idx][0].record(stream)
        net.infer(q_d, out_d, nq, sp)
        if step_idx is not None:
            ev_inf[step_idx][1].record(stream)
        if world == 1 and not args.per_batch_launch:
            net.train_batches(q_ptr_t, t_ptr_t, NUM_BATCHES, sp)     # Device::nrcTrainRadiance: 4 steps, one launch
        else:
            for b in range(NUM_BATCHES):
                train_batch(b)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(args.warmup, 3)):
        update()
    barrier()
    sampler = ClockSampler(local)
    sampler.start()
    launches0 = net.launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record(stream)
    for i in range(args.steps):
        update(i)
    e1.record(stream)
    sampler.sample_now()                     # the launches above are still executing: at least one sample under load
    barrier()
    clocks = sampler.stop()
    launches = net.launch_count() - launches0
    ms_total = e0.elapsed_time(e1)
    t_inf_ms = float(np.mean([a.elapsed_time(b) for a, b in ev_inf]))
    if world > 1:
        t = torch.tensor([ms_total], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_total = float(t.item())
    ms_step = ms_total / args.steps
    total_units = (args.shard_frame + n_train) if strong else world * (nq + n_train)
    value = total_units / (ms_step * 1e-3)

    # ---- e2e: host buffers through the C ABI (H2D of queries / batches and D2H of radiance + loss inside)
    def update_host():
        net.infer_host(q_h, out_h, nq)
        loss = None
        for b in range(NUM_BATCHES):
            qs = tq_h.data_ptr() + b * BATCH_SIZE * 60
            ts = tt_h.data_ptr() + b * BATCH_SIZE * 12
            if world == 1:
                loss = net.train_host(qs, ts, want_loss=(b == NUM_BATCHES - 1))
            else:   # DP: stage the batch, gradient, all-reduce, apply
                tq_d[b * BATCH_SIZE:(b + 1) * BATCH_SIZE].copy_(tq_h[b * BATCH_SIZE:(b + 1) * BATCH_SIZE], non_blocking=True)
                tt_d[b * BATCH_SIZE:(b + 1) * BATCH_SIZE].copy_(tt_h[b * BATCH_SIZE:(b + 1) * BATCH_SIZE], non_blocking=True)
                loss = train_batch(b, want_loss=(b == NUM_BATCHES - 1))
        return loss

    e2e_steps = max(1, min(args.e2e_steps, args.steps))
    for _ in range(2):
        update_host()
    barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        update_host()
    barrier()
    e2e_s = (time.perf_counter() - t0) / e2e_steps
    if world > 1:
        t = torch.tensor([e2e_s], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_s = float(t.item())
    e2e = {"value": total_units / e2e_s, "unit": UNIT, "ms_per_step": e2e_s * 1e3, "steps": e2e_steps,
           "h2d_bytes_per_step": nq * 60 + n_train * 72, "d2h_bytes_per_step": nq * 12 + 4,
           "timing": "host wall clock between device synchronisations, max over ranks"}

    # ---- the whole NRC sequence of Device::render through the frame driver (NEXT rows included): infer -> accumulate ->
    #      propagate -> fused shuffle -> 4 x train, device-resident buffers, CUDA events
    frame_ms = None
    if world == 1:
        n_px, n_tiles = 2560 * 1440, FRAME_QUERIES - 2560 * 1440
        if nq == FRAME_QUERIES:
            rng = np.random.default_rng(5)
            thr_d = torch.from_numpy(rng.uniform(0, 1, (n_px, 3)).astype(np.float32)).cuda()
            out4_d = torch.zeros((n_px, 4), device="cuda")
            ev_np, rec_np, used = synth.training_chains(n_tiles)
            ev_d, rec_d = torch.from_numpy(ev_np).cuda(), torch.from_numpy(rec_np).cuda()
            tgt_work = tt_d.clone()
            fr = pkg.Frame(q_d.data_ptr(), out_d.data_ptr(), n_px, n_tiles, thr_d.data_ptr(), out4_d.data_ptr(), pkg.RENDER_FULL, 0,
                           ev_d.data_ptr(), rec_d.data_ptr(), tq_d.data_ptr(), tgt_work.data_ptr(), used, 0)
            for i in range(3):
                net.update(fr, sp)
            torch.cuda.synchronize()
            f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            n_fr = max(10, min(args.steps, 100))
            f0.record(stream)
            for i in range(n_fr):
                fr.iteration_index, fr.shuffle_seed = i, i
                tgt_work.copy_(tt_d)              # Device::render zeroes / refills the targets every frame (Device.cpp:2471-2476)
                net.update(fr, sp)
            f1.record(stream)
            torch.cuda.synchronize()
            frame_ms = f0.elapsed_time(f1) / n_fr

    # ---- the OTHER encoding on the same buffers (BASELINE.json configs[2] names the 2K frame with the HashGrid encoding;
    #      the headline line is quoted on the reference's default, Frequency): a short device-timed run, reported as extra keys
    other = None
    if world == 1 and not args.no_other_encoding:
        oenc = 1 - enc
        onet = pkg.Network()
        onet.init(sp, oenc)
        def oupdate():
            onet.infer(q_d, out_d, nq, sp)
            onet.train_batches(q_ptr_t, t_ptr_t, NUM_BATCHES, sp)
        for _ in range(3):
            oupdate()
        torch.cuda.synchronize()
        n_o = max(10, min(args.steps, 50))
        o0, o1, o2 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
        t_oi = 0.0
        o0.record(stream)
        for _ in range(n_o):
            oupdate()
        o1.record(stream)
        for _ in range(n_o):
            onet.infer(q_d, out_d, nq, sp)
        o2.record(stream)
        torch.cuda.synchronize()
        o_ms, o_inf = o0.elapsed_time(o1) / n_o, o1.elapsed_time(o2) / n_o
        other = {"workload": workload_name(oenc, nq), "ms_per_step": o_ms, "ms_infer": o_inf, "ms_train_4_steps": o_ms - o_inf,
                 "value": (nq + n_train) / (o_ms * 1e-3), "unit": UNIT, "steps": n_o}
        onet.destroy()

    # ---- roofline of the dominant kernel (fused query kernel), live CUDA-event duration per launch
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak_tf = peaks.get("bf16_tflops", 1590.0)
    achieved_tf = FLOP_PER_QUERY[enc] * nq / (t_inf_ms * 1e-3) / 1e12
    roofline = {"kernel": "nrc_infer_kernel", "bound": "tensor", "achieved": achieved_tf, "peak": peak_tf,
                "unit": "TFLOP/s", "frac": achieved_tf / peak_tf,
                "peak_source": "MEASURED_PEAKS.json bf16_tflops (burst), of measured" if peaks else "fallback 1590, of fallback",
                "frac_of_sustained": achieved_tf / peaks["bf16_tflops_sustained"] if "bf16_tflops_sustained" in peaks else None,
                "us_per_launch": t_inf_ms * 1e3, "share_of_step": t_inf_ms / ms_step,
                "hbm_gbs_algorithmic": 72.0 * nq / (t_inf_ms * 1e-3) / 1e9,
                "traffic": TRAFFIC_BYTES.get((enc, nq))}

    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
            "ms_per_step": ms_step, "higher_is_better": True, "scaling": "strong" if strong else "weak", "vs_baseline": None, "dtype": "f16",
            "data": "synthetic",
            "config": {"workload": (f"ONE frame of {args.shard_frame} queries batch-sharded over {world} GPUs + {NUM_BATCHES}x{BATCH_SIZE} "
                                    f"training samples split across them, {'Frequency' if enc == 0 else 'Hash'} encoding, 5x64 ReLU MLP")
                                   if strong else workload_name(enc, nq),
                       "queries_per_gpu": nq, "train_samples_per_gpu": NUM_BATCHES * train_local if world > 1 else n_train,
                       "parallelism": (f"dp{world} ({args.dp} all-reduce)") if world > 1 else "single",
                       "query_order": ("pixel-ordered primary hits of a pinhole camera in the Cornell box + incoherent "
                                       "tile tails" if args.query_order == "frame" else
                                       "independent random surface points (incoherent worst case for the hash grid)"),
                       "l2": f"inputs larger than L2 ({nq * 60 / 1e6:.0f} MB of queries per step vs 126 MB L2)"},
            "queries_per_s": (args.shard_frame if strong else world * nq) / (t_inf_ms * 1e-3),
            "train_samples_per_s": (n_train if strong else world * n_train) / max((ms_step - t_inf_ms) * 1e-3, 1e-9),
            "ms_infer": t_inf_ms, "ms_train_4_steps": ms_step - t_inf_ms,
            "frame_driver_ms": frame_ms, "other_encoding": other, "clocks": clocks, "e2e": e2e, "gpu_launches": int(launches), "roofline": roofline}
    net.destroy()
    return line, rank, world


def time_reference_gpu(enc, nq):
    """The reference's own CUDA build on this GPU (subprocess; test infrastructure under oracle/)."""
    import subprocess
    so = os.path.join(ROOT, "oracle", "_ref", "libnrc_ref.so")
    if not os.path.exists(so):
        return {"unavailable": "oracle/_ref/libnrc_ref.so not built"}
    r = subprocess.run([sys.executable, os.path.join(ROOT, "oracle", "ref", "time_ref_gpu.py"), str(enc), str(nq)],
                       capture_output=True, text=True, timeout=600)
    if r.returncode != 0:
        return {"unavailable": "reference run failed: " + (r.stderr or r.stdout)[-300:]}
    return json.loads(r.stdout.strip().splitlines()[-1])


# (encoding, queries) -> dram__bytes_read.sum + dram__bytes_write.sum of ONE launch of the query kernel, from the
# `ncu --set full` captures summarised in profiles/r01_infer_{frequency,hash}_ncu.txt (algorithmic: 72 B/query = 269.6 MB)
TRAFFIC_BYTES = {(0, FRAME_QUERIES): 224733696 + 31894272, (1, FRAME_QUERIES): 226329856 + 33507840}


def main():
    # Libraries (NCCL version banner ...) write to fd 1; the contract is ONE JSON line on stdout, so fd 1 is pointed at
    # stderr for the duration of the run and the result goes to the saved descriptor.
    sys.stdout.flush()
    real_stdout = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    _main(real_stdout)


def _main(out):
    args = parse_args()
    enc = 0 if args.encoding == "frequency" else 1
    if args.impl == "reference":
        run_reference(args, enc, out)
        return
    line, rank, world = run_ours(args, enc)
    if rank == 0:
        if world == 1 and not args.no_cpu_baseline:
            orc = entry.load_oracle()
            synth = entry.load_package_synthetic()
            threads = orc.ncpu()
            v, t_update, sample = cpu_update_rate(orc, synth, enc, args.queries, threads)
            line["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample,
                                    "ms_per_update": t_update * 1e3}
        if args.ref_gpu and world == 1:
            line["reference_gpu"] = time_reference_gpu(enc, args.queries)
        print(json.dumps(line), file=out, flush=True)
    if world > 1:
        import torch.distributed as dist
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
