#!/usr/bin/env python
"""bench.py -- the NRC update benchmark (BASELINE.json metric) on N B200s of one node.

A "step" is ONE NRC UPDATE of a 2K frame, the per-frame hot path of the reference's Device::render
(/root/reference/nrc/src/Device.cpp:2466-2515):
    Network::infer on (#px + #tiles) = 2560*1440 + 320*180 = 3 744 000 RadianceQuery records, then
    NUM_BATCHES = 4 sequential Network::train steps of BATCH_SIZE = 16384 path-tail samples each.
Headline line: Frequency encoding (the reference's default, Device.cpp:274) unless --encoding hash.  The same JSON line
carries, under "configs", a FULL record (ms_per_step, e2e, rooflines, reference build on this GPU) of the configuration
BASELINE.json configs[2] names -- the 2K frame with the HashGrid encoding -- and the 2^20-query case of configs[0].

value        = (queries + training samples) processed per second over all ranks, inputs resident in HBM.
e2e          = the same through the host-buffer C-ABI calls (pinned host memory in, results back to host).
roofline     = the dominant kernel (fused query kernel) against the measured tensor peak; "rooflines" lists every kernel.
cpu_baseline = the scalar CPU oracle on all host threads (the reference has no CPU path; oracle/ is the port), one REAL
               full update.  --impl reference runs that for every step.
reference_gpu = the reference's own tcnn build for sm_100 (oracle/_ref/libnrc_ref.so) on this GPU, when it was built.

Multi-GPU (torchrun): every rank owns a full 2K frame of queries (batch-sharded inference, no collective) and 16384 local
training samples per step; the fp32 weight gradients are summed INSIDE the step kernel over NVLink peer memory (default
--dp fused; the frame's four steps are one launch per rank) or with an NCCL all-reduce between the backward kernel and the
Adam+EMA kernel (--dp nccl); global batch = N*16384, so per-GPU work is fixed: "scaling": "weak".  The line also carries
"strong_4k" (BASELINE.json configs[3]): ONE 4K frame batch-sharded over the ranks, training batches split across them.
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
FRAME_4K_QUERIES = 3840 * 2160 + (3840 // 8) * (2160 // 8)   # 8 424 000
NUM_BATCHES, BATCH_SIZE = 4, 16384
FLOP_PER_QUERY = {0: 45056, 1: 43008}                          # 2*P_mlp
FLOP_PER_TRAIN_SAMPLE = {0: 124928, 1: 129024}
ENC_NAME = {0: "Frequency", 1: "Hash"}


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", choices=["ours", "reference"], default="ours")
    ap.add_argument("--encoding", choices=["frequency", "hash"], default="frequency")
    ap.add_argument("--queries", type=int, default=FRAME_QUERIES, help="queries per update per GPU")
    ap.add_argument("--query-order", choices=["frame", "random"], default="frame",
                    help="frame: pixel-ordered primary hits of a camera in the Cornell box + incoherent tile tails (what a "
                         "renderer hands over); random: independent random surface points (incoherent worst case)")
    ap.add_argument("--per-batch-launch", action="store_true",
                    help="one nrc_b200_train launch per batch (the reference's call pattern) instead of nrc_b200_train_batches")
    ap.add_argument("--no-configs", action="store_true", help="skip the extra records (other encoding, 2^20 queries, strong 4K)")
    ap.add_argument("--shard-frame", type=int, default=0, metavar="TOTAL",
                    help="strong scaling as the HEADLINE (BASELINE.json configs[3]): ONE frame of TOTAL queries (4K: 8424000) "
                         "batch-sharded over the ranks, 16384-sample training batches split across the ranks")
    ap.add_argument("--e2e-steps", type=int, default=10)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--dp", choices=["fused", "nccl"], default="fused",
                    help="multi-GPU gradient all-reduce: fused into the step kernel over NVLink peer memory, or NCCL")
    ap.add_argument("--no-ref-gpu", action="store_true",
                    help="do not time the reference's own tcnn build (oracle/_ref/libnrc_ref.so) on this GPU")
    return ap.parse_args()


def workload_name(enc, nq):
    return (f"2K-frame NRC update: infer {nq} queries (2560x1440 px + 320x180 tiles) + {NUM_BATCHES}x{BATCH_SIZE} "
            f"training samples, {ENC_NAME[enc]} encoding, 5x64 ReLU MLP")


def config_dict(enc, nq, world, dp="fused", order="frame", strong_total=0):
    """The `config` object of BOTH arms (ours and --impl reference print the identical dict for the same flags)."""
    wl = (f"ONE frame of {strong_total} queries batch-sharded over {world} GPUs + {NUM_BATCHES}x{BATCH_SIZE} training samples "
          f"split across them, {ENC_NAME[enc]} encoding, 5x64 ReLU MLP") if strong_total else workload_name(enc, nq)
    return {"workload": wl, "queries_per_gpu": nq,
            "train_samples_per_gpu": NUM_BATCHES * BATCH_SIZE // (world if strong_total else 1),
            "parallelism": f"dp{world} ({dp} all-reduce)" if world > 1 else "single",
            "query_order": ("pixel-ordered primary hits of a pinhole camera in the Cornell box + incoherent tile tails"
                            if order == "frame" else "independent random surface points (incoherent worst case for the hash grid)"),
            "l2": f"inputs larger than L2 ({nq * 60 / 1e6:.0f} MB of queries per step vs 126 MB L2)"}


def make_inputs(synth, nq, rank, order="frame", width=2560):
    """Seeded Cornell-box queries (host, float32).  order == "frame": `width`-pixel-wide rows of primary hits in pixel order,
    the last 1/65 being incoherent tile tails (8x8 tiles); "random": independent surface points, tiled from 2^20-query blocks
    to keep start-up short."""
    if order == "frame":
        n_tail = 320 * 180 if nq == FRAME_QUERIES else nq // 65
        rows = -(-(nq - n_tail) // width)
        px = synth.cornell_frame(width, rows, seed=1337 + 97 * rank)[:nq - n_tail]
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


def bind_to_gpu_numa(local):
    """Pins this process (and therefore the first-touch placement of its pinned host buffers) to the CPUs of the NUMA node
    the GPU hangs off: with 8 ranks each pushing 274 MB per update through PCIe, buffers that all live on one socket make
    the host memory system, not the kernels, the e2e limit.  Best effort; returns what it did."""
    try:
        import torch
        bus = torch.cuda.get_device_properties(local).pci_bus_id
        dom = torch.cuda.get_device_properties(local).pci_domain_id
        dev = torch.cuda.get_device_properties(local).pci_device_id
        path = f"/sys/bus/pci/devices/{dom:04x}:{bus:02x}:{dev:02x}.0/numa_node"
        node = int(open(path).read().strip())
        if node < 0:
            return {"numa_node": node, "bound": False}
        cpus = []
        for part in open(f"/sys/devices/system/node/node{node}/cpulist").read().strip().split(","):
            a, _, b = part.partition("-")
            cpus += list(range(int(a), int(b or a) + 1))
        allowed = sorted(set(cpus) & set(os.sched_getaffinity(0)))
        if allowed:
            os.sched_setaffinity(0, allowed)
        return {"numa_node": node, "bound": bool(allowed), "cpus": len(allowed)}
    except Exception as e:
        return {"numa_node": None, "bound": False, "error": repr(e)[:120]}


# --------------------------------------------------------------------------------------------- CPU arms (oracle port)
def cpu_full_update(orc, tr, q, tq, tt, threads):
    """One REAL NRC update on the CPU oracle: infer on all of q, then NUM_BATCHES optimizer steps.  Returns seconds."""
    t0 = time.perf_counter()
    orc.infer(tr.spec, tr.ema, q, nthreads=threads)
    for b in range(NUM_BATCHES):
        tr.train_step(tq[b * BATCH_SIZE:(b + 1) * BATCH_SIZE], tt[b * BATCH_SIZE:(b + 1) * BATCH_SIZE], nthreads=threads)
    return time.perf_counter() - t0


def cpu_probe(orc, tr, synth, threads):
    """Per-unit CPU times from a small probe (us per query, us per training sample)."""
    q = synth.cornell_queries(8192, seed=6)
    t0 = time.perf_counter()
    orc.infer(tr.spec, tr.ema, q, nthreads=threads)
    t_q = (time.perf_counter() - t0) / 8192
    tq = synth.cornell_queries(2048, seed=8)
    tt = synth.radiance_targets(tq, seed=8)
    t0 = time.perf_counter()
    tr.train_step(tq, tt, nthreads=threads)
    t_s = (time.perf_counter() - t0) / 2048
    return t_q, t_s


def cpu_baseline_leg(orc, synth, enc, q_np, tq_np, tt_np, threads, budget_s=45.0):
    """cpu_baseline of our arm: ONE real full update on all host threads when the probe says it fits the budget, else a
    same-mix sample (57.13 queries per training sample) whose units/time is reported."""
    tr = orc.Trainer(enc)
    tr.ema[:] = tr.params                      # query with live weights (EMA starts at zero)
    t_q, t_s = cpu_probe(orc, tr, synth, threads)
    nq, n_train = q_np.shape[0], NUM_BATCHES * BATCH_SIZE
    est = nq * t_q + n_train * t_s
    if est <= budget_s:
        dt = cpu_full_update(orc, tr, q_np, tq_np, tt_np, threads)
        units, sample = nq + n_train, f"one REAL full update ({nq} queries + {n_train} training samples) on {threads} host threads, {dt:.2f} s"
    else:
        f = budget_s / est
        n_tr = max(128, int(BATCH_SIZE * f) // 128 * 128)
        n_inf = int(nq * n_tr / BATCH_SIZE)
        t0 = time.perf_counter()
        orc.infer(tr.spec, tr.ema, q_np[:n_inf], nthreads=threads)
        for b in range(NUM_BATCHES):
            tr.train_step(tq_np[b * BATCH_SIZE:b * BATCH_SIZE + n_tr], tt_np[b * BATCH_SIZE:b * BATCH_SIZE + n_tr], nthreads=threads)
        dt = time.perf_counter() - t0
        units = n_inf + NUM_BATCHES * n_tr
        sample = (f"same-mix sample of the update: {n_inf} queries + {NUM_BATCHES}x{n_tr} training samples on {threads} host threads, "
                  f"{dt:.2f} s (a full update is estimated at {est:.1f} s)")
    tr.close()
    return {"value": units / dt, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample,
            "ms_per_update": (nq + n_train) / (units / dt) * 1e3,
            "us_per_query_probe": t_q * 1e6, "us_per_train_sample_probe": t_s * 1e6}


def run_reference(args, enc, out):
    """--impl reference: the reference has no CPU implementation of this path (every function is __global__); the timed
    code is oracle/'s scalar C restatement (kind "port") on all host threads.  Every step EXECUTES the whole update
    (3 744 000 queries + 4 x 16384 training samples) when warmup + steps of it fit ~10 minutes (the driver's 20 + 5 steps
    do: ~3.5 s each on 16 threads); otherwise every step is a same-mix sample and `value` is that sample's units / time."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    orc = entry.load_oracle()
    orc.build()
    synth = entry.load_package_synthetic()
    threads = orc.ncpu()
    tr = orc.Trainer(enc)
    tr.ema[:] = tr.params
    t_q, t_s = cpu_probe(orc, tr, synth, threads)
    nq, n_train = args.queries, NUM_BATCHES * BATCH_SIZE
    est = nq * t_q + n_train * t_s
    n_iter = args.steps + args.warmup
    full = est * n_iter <= 600.0
    if full:
        q, tq, tt = make_inputs(synth, nq, 0, args.query_order)
        n_inf, n_tr = nq, BATCH_SIZE
    else:
        f = min(1.0, 240.0 / (est * n_iter))
        n_tr = max(128, int(BATCH_SIZE * f) // 128 * 128)
        n_inf = max(128, int(nq * n_tr / BATCH_SIZE))
        q, tq, tt = make_inputs(synth, n_inf, 0, "random")
    times = []
    for it in range(n_iter):
        t0 = time.perf_counter()
        orc.infer(tr.spec, tr.ema, q[:n_inf], nthreads=threads)
        for b in range(NUM_BATCHES):
            tr.train_step(tq[b * BATCH_SIZE:b * BATCH_SIZE + n_tr], tt[b * BATCH_SIZE:b * BATCH_SIZE + n_tr], nthreads=threads)
        if it >= args.warmup:
            times.append(time.perf_counter() - t0)
    tr.close()
    units = n_inf + NUM_BATCHES * n_tr
    dt = float(np.mean(times))
    v = units / dt
    sample = (f"every step is one REAL full update: {n_inf} queries + {NUM_BATCHES}x{n_tr} training samples on {threads} host threads"
              if full else
              f"every step is a same-mix sample of the update ({n_inf} queries + {NUM_BATCHES}x{n_tr} training samples on {threads} host "
              f"threads; a full update is estimated at {est:.1f} s, {n_iter} of them would not fit); value = units of the sample / its time")
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f16", "data": "synthetic",
            "config": config_dict(enc, args.queries, max(1, args.gpus), args.dp, args.query_order),
            "cpu_baseline": {"value": v, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample,
                             "ms_per_update": (nq + n_train) / v * 1e3, "executed_full_update": bool(full)},
            "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), file=out, flush=True)


# --------------------------------------------------------------------------------------------- the GPU arm
def load_peaks():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        return {}


def load_traffic():
    """dram__bytes_read.sum + dram__bytes_write.sum per launch from the committed `ncu --set full` summaries
    (profiles/r02_traffic.json, written by tools/ncu_summary.py --traffic): {key: {"bytes": ..., "source": file}}."""
    for name in ("r02_traffic.json", "r01_traffic.json"):
        p = os.path.join(ROOT, "profiles", name)
        if os.path.exists(p):
            try:
                return json.load(open(p))
            except Exception:
                pass
    return {}


def measure_l2_peak(pkg):
    """Measured L2 read bandwidth of this GPU: the library's probe kernel (16-byte L1-bypassing loads over a 32 MiB buffer that
    stays resident in the 126 MB L2, best of 5 x 20 passes) -- the denominator of the hash-grid gather roofline."""
    return float(pkg.l2_read_bandwidth(32 << 20, 20))


def tensor_roofline(kernel, cfg, flop, us, peaks, traffic=None, extra=None):
    peak = peaks.get("bf16_tflops", 1590.0)
    ach = flop / (us * 1e-6) / 1e12
    r = {"kernel": kernel, "config": cfg, "bound": "tensor", "achieved": ach, "peak": peak, "unit": "TFLOP/s", "frac": ach / peak,
         "peak_source": "MEASURED_PEAKS.json bf16_tflops (burst)" if peaks else "fallback 1590 (B200_PROFILING.md)",
         "frac_of_sustained": ach / peaks["bf16_tflops_sustained"] if "bf16_tflops_sustained" in peaks else None,
         "us_per_launch": us, "traffic": traffic["bytes"] if traffic else None,
         "traffic_source": traffic["source"] if traffic else None}
    if extra:
        r.update(extra)
    return r


def time_reference_gpu(enc, nq):
    """The reference's own CUDA build on this GPU (subprocess; test infrastructure under oracle/)."""
    import subprocess
    so = os.path.join(ROOT, "oracle", "_ref", "libnrc_ref.so")
    if not os.path.exists(so):
        return {"unavailable": "oracle/_ref/libnrc_ref.so not built"}
    try:
        r = subprocess.run([sys.executable, os.path.join(ROOT, "oracle", "ref", "time_ref_gpu.py"), str(enc), str(nq)],
                           capture_output=True, text=True, timeout=600)
    except Exception as e:
        return {"unavailable": "reference run failed: " + repr(e)[:200]}
    if r.returncode != 0:
        return {"unavailable": "reference run failed: " + (r.stderr or r.stdout)[-300:]}
    return json.loads(r.stdout.strip().splitlines()[-1])


class Bench:
    """One network + its device / pinned buffers; measures the update of one configuration."""

    def __init__(self, torch, pkg, synth, enc, nq, rank, world, args, width=2560, train_local=BATCH_SIZE, dp=None, strong=False):
        self.torch, self.pkg, self.enc, self.nq, self.rank, self.world, self.args = torch, pkg, enc, nq, rank, world, args
        self.train_local, self.strong = train_local, strong
        self.q_np, self.tq_np, self.tt_np = make_inputs(synth, nq, rank, args.query_order, width)
        self.q_h = torch.from_numpy(self.q_np).pin_memory()
        self.tq_h = torch.from_numpy(self.tq_np).pin_memory()
        self.tt_h = torch.from_numpy(self.tt_np).pin_memory()
        self.out_h = torch.empty((nq, 3), dtype=torch.float32).pin_memory()
        self.q_d, self.tq_d, self.tt_d = self.q_h.cuda(), self.tq_h.cuda(), self.tt_h.cuda()
        self.out_d = torch.empty((nq, 3), device="cuda")
        self.stream = torch.cuda.current_stream()
        self.sp = self.stream.cuda_stream
        self.net = pkg.Network()
        self.net.init(self.sp, enc)
        self.dp = dp(self.net) if dp else None
        self.q_ptr, self.t_ptr = self.tq_d.data_ptr(), self.tt_d.data_ptr()

    def train(self):
        net, sp, a = self.net, self.sp, self.args
        if self.world == 1:
            if a.per_batch_launch:
                for b in range(NUM_BATCHES):
                    net.train(self.q_ptr + b * BATCH_SIZE * 60, self.t_ptr + b * BATCH_SIZE * 12, sp)
            else:
                net.train_batches(self.q_ptr, self.t_ptr, NUM_BATCHES, sp)     # Device::nrcTrainRadiance: 4 steps, one launch
        elif a.dp == "fused" and not self.strong and not a.per_batch_launch:
            net.train_batches(self.q_ptr, self.t_ptr, NUM_BATCHES, sp)         # connected handle: global batch, all-reduce inside
        else:
            off = self.rank * self.train_local if self.strong else 0           # strong: this rank's slice of the SAME batch
            for b in range(NUM_BATCHES):
                self.dp.train(self.q_ptr + (b * BATCH_SIZE + off) * 60, self.t_ptr + (b * BATCH_SIZE + off) * 12, self.train_local, sp)

    def barrier(self):
        if self.world > 1:
            import torch.distributed as dist
            dist.barrier()
        self.torch.cuda.synchronize()

    def timed(self, steps, warmup, sampler=None):
        """W warm-up updates, then EXACTLY `steps` timed ones bracketed by barrier + synchronize; returns
        (ms per update, ms infer, ms train, launches), max over ranks for the update."""
        torch, net = self.torch, self.net
        ev = [[torch.cuda.Event(enable_timing=True) for _ in range(3)] for _ in range(steps)]
        for _ in range(max(warmup, 3)):
            net.infer(self.q_d, self.out_d, self.nq, self.sp)
            self.train()
        self.barrier()
        if sampler:
            sampler.start()
        l0 = net.launch_count()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        self.barrier()
        e0.record(self.stream)
        for i in range(steps):
            ev[i][0].record(self.stream)
            net.infer(self.q_d, self.out_d, self.nq, self.sp)
            ev[i][1].record(self.stream)
            self.train()
            ev[i][2].record(self.stream)
        e1.record(self.stream)
        if sampler:
            sampler.sample_now()             # the launches above are still executing: at least one sample under load
        self.barrier()
        launches = net.launch_count() - l0
        ms_total = e0.elapsed_time(e1)
        t_inf = float(np.mean([a.elapsed_time(b) for a, b, _ in ev]))
        t_tr = float(np.mean([b.elapsed_time(c) for _, b, c in ev]))
        if self.world > 1:
            import torch.distributed as dist
            t = torch.tensor([ms_total], device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms_total = float(t.item())
        return ms_total / steps, t_inf, t_tr, int(launches)

    def e2e(self, steps):
        """The same update through the host-buffer C-ABI calls: every step copies its inputs from pinned host memory and
        reads the radiance and the last loss back; host wall clock between device synchronisations, max over ranks."""
        torch, net = self.torch, self.net

        def update_host():
            net.infer_host(self.q_h, self.out_h, self.nq)
            loss = None
            for b in range(NUM_BATCHES):
                qs = self.tq_h.data_ptr() + b * BATCH_SIZE * 60
                ts = self.tt_h.data_ptr() + b * BATCH_SIZE * 12
                last = b == NUM_BATCHES - 1
                if self.world == 1:
                    loss = net.train_host(qs, ts, want_loss=last)
                else:   # connected handle / DP trainer: stage the batch, then the data-parallel step
                    sl = slice(b * BATCH_SIZE, (b + 1) * BATCH_SIZE)
                    self.tq_d[sl].copy_(self.tq_h[sl], non_blocking=True)
                    self.tt_d[sl].copy_(self.tt_h[sl], non_blocking=True)
                    off = self.rank * self.train_local if self.strong else 0
                    loss = self.dp.train(self.q_ptr + (b * BATCH_SIZE + off) * 60, self.t_ptr + (b * BATCH_SIZE + off) * 12,
                                         self.train_local, self.sp, want_loss=last)
            return loss

        for _ in range(2):
            update_host()
        self.barrier()
        t0 = time.perf_counter()
        for _ in range(steps):
            update_host()
        self.barrier()
        s = (time.perf_counter() - t0) / steps
        if self.world > 1:
            import torch.distributed as dist
            t = torch.tensor([s], device="cuda", dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            s = float(t.item())
        n_train = NUM_BATCHES * BATCH_SIZE
        h2d, d2h = self.nq * 60 + n_train * 72, self.nq * 12 + 4
        return s, {"ms_per_step": s * 1e3, "steps": steps, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                   "pcie_gbs_per_rank": (h2d + d2h) / s / 1e9,
                   "timing": "host wall clock between device synchronisations, max over ranks"}

    def infer_flushed(self, n, reps=20):
        """Per-launch time of the query kernel on the first n queries with an L2 flush (256 MB memset) between launches --
        for batches that fit in L2 (BASELINE.json configs[0]: 2^20 queries = 63 MB)."""
        torch = self.torch
        flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
        ts = []
        for i in range(reps + 3):
            flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(self.stream)
            self.net.infer(self.q_d, self.out_d, n, self.sp)
            e1.record(self.stream)
            torch.cuda.synchronize()
            if i >= 3:
                ts.append(e0.elapsed_time(e1) * 1e3)
        return float(np.median(ts)), float(np.min(ts))

    def close(self):
        self.net.destroy()


def rooflines_for(enc, nq, t_inf_ms, t_tr_ms, peaks, traffic, l2_peak):
    """The kernels of one configuration against what bounds them."""
    key = f"infer_{ENC_NAME[enc].lower()}_{nq}"
    rl = [tensor_roofline(f"nrc_infer_kernel<{enc}>", f"{nq} queries, {ENC_NAME[enc]}", FLOP_PER_QUERY[enc] * nq, t_inf_ms * 1e3, peaks,
                          traffic.get(key), {"hbm_gbs_algorithmic": 72.0 * nq / (t_inf_ms * 1e-3) / 1e9})]
    if enc == 1:     # the hash-grid gathers: 16 levels x 8 corners x 4 B per query through L1/L2 (table 1.98 MB << L2)
        ach = 512.0 * nq / (t_inf_ms * 1e-3) / 1e9
        t = traffic.get(key + "_l2")
        rl.append({"kernel": "nrc_infer_kernel<1> (hash-grid gathers)", "config": f"{nq} queries, Hash", "bound": "l2", "achieved": ach,
                   "peak": l2_peak, "unit": "GB/s", "frac": ach / l2_peak,
                   "peak_source": "L2-resident 32 MiB read probe measured in this run (nrc_b200_debug_l2_read_bandwidth)",
                   "us_per_launch": t_inf_ms * 1e3, "traffic": t["bytes"] if t else None, "traffic_source": t["source"] if t else None})
    n_tr = NUM_BATCHES * BATCH_SIZE
    rl.append(tensor_roofline(f"nrc_train_step_kernel<{enc}>", f"{NUM_BATCHES} x {BATCH_SIZE} samples (BASELINE configs[1]), {ENC_NAME[enc]}",
                              FLOP_PER_TRAIN_SAMPLE[enc] * n_tr, t_tr_ms * 1e3, peaks, traffic.get(f"train_{ENC_NAME[enc].lower()}"),
                              {"us_per_optimizer_step": t_tr_ms * 1e3 / NUM_BATCHES,
                               "note": "latency-bound: 4 dependent optimizer steps of 11 dependent MMA->epilogue trips each; tensor floor 4.9 us"}))
    return rl


def run_ours(args, enc):
    import torch
    import torch.distributed as dist
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a B200: there is no CPU path (use --impl reference for the CPU oracle)")
    torch.cuda.set_device(local)
    numa = bind_to_gpu_numa(local) if world > 1 else None
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    pkg = entry.load_package()
    synth = entry.load_package_synthetic()
    distmod = importlib.import_module("nrc_b200.dist")
    n_train = NUM_BATCHES * BATCH_SIZE
    strong = args.shard_frame > 0
    nq, train_local = args.queries, BATCH_SIZE
    if strong:                                   # one frame for the whole job: contiguous tile-aligned shard per rank
        lo, hi = distmod.shard_range(args.shard_frame, rank, world) if world > 1 else (0, args.shard_frame)
        nq, train_local = hi - lo, BATCH_SIZE // world
    dp_factory = (lambda net: distmod.DataParallelTrainer(net, mode=args.dp)) if world > 1 else None
    peaks, traffic = load_peaks(), load_traffic()

    B = Bench(torch, pkg, synth, enc, nq, rank, world, args, train_local=train_local, dp=dp_factory, strong=strong)
    sampler = ClockSampler(local)
    ms_step, t_inf, t_tr, launches = B.timed(args.steps, args.warmup, sampler)
    clocks = sampler.stop()
    total_units = (args.shard_frame + n_train) if strong else world * (nq + n_train)
    value = total_units / (ms_step * 1e-3)
    e2e_s, e2e = B.e2e(max(1, min(args.e2e_steps, args.steps)))
    e2e = dict({"value": total_units / e2e_s, "unit": UNIT}, **e2e)
    if numa:
        e2e["numa"] = numa
    l2_peak = measure_l2_peak(pkg) if rank == 0 else 1.0
    rl = rooflines_for(enc, nq, t_inf, t_tr, peaks, traffic, l2_peak)
    roofline = dict(rl[0], share_of_step=t_inf / ms_step)

    # ---- the whole NRC sequence of Device::render through the frame driver (NEXT rows included): infer -> accumulate ->
    #      propagate -> fused shuffle -> 4 x train, device-resident buffers, CUDA events
    frame_ms = None
    if world == 1 and nq == FRAME_QUERIES:
        n_px, n_tiles = 2560 * 1440, FRAME_QUERIES - 2560 * 1440
        rng = np.random.default_rng(5)
        thr_d = torch.from_numpy(rng.uniform(0, 1, (n_px, 3)).astype(np.float32)).cuda()
        out4_d = torch.zeros((n_px, 4), device="cuda")
        ev_np, rec_np, used = synth.training_chains(n_tiles)
        ev_d, rec_d = torch.from_numpy(ev_np).cuda(), torch.from_numpy(rec_np).cuda()
        tgt_work = B.tt_d.clone()
        fr = pkg.Frame(B.q_d.data_ptr(), B.out_d.data_ptr(), n_px, n_tiles, thr_d.data_ptr(), out4_d.data_ptr(), pkg.RENDER_FULL, 0,
                       ev_d.data_ptr(), rec_d.data_ptr(), B.tq_d.data_ptr(), tgt_work.data_ptr(), used, 0)
        for i in range(3):
            B.net.update(fr, B.sp)
        torch.cuda.synchronize()
        f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        n_fr = max(10, min(args.steps, 100))
        f0.record(B.stream)
        for i in range(n_fr):
            fr.iteration_index, fr.shuffle_seed = i, i
            tgt_work.copy_(B.tt_d)              # Device::render zeroes / refills the targets every frame (Device.cpp:2471-2476)
            B.net.update(fr, B.sp)
        f1.record(B.stream)
        torch.cuda.synchronize()
        frame_ms = f0.elapsed_time(f1) / n_fr
        del thr_d, out4_d

    configs = []
    extra_rl = []
    if world == 1 and not args.no_configs and nq == FRAME_QUERIES:
        # ---- BASELINE.json configs[0]: 2^20 queries (fits L2: flushed between launches), same network
        med, mn = B.infer_flushed(1 << 20)
        r0 = tensor_roofline(f"nrc_infer_kernel<{enc}>", f"BASELINE configs[0]: 2^20 queries, {ENC_NAME[enc]}, L2 flushed between launches",
                             FLOP_PER_QUERY[enc] * (1 << 20), med, peaks, traffic.get(f"infer_{ENC_NAME[enc].lower()}_{1 << 20}"),
                             {"us_min": mn, "queries_per_s": (1 << 20) / (med * 1e-6)})
        extra_rl.append(r0)
    B.close()
    q_np, tq_np, tt_np = B.q_np, B.tq_np, B.tt_np
    del B
    torch.cuda.empty_cache()

    if world == 1 and not args.no_configs and nq == FRAME_QUERIES:
        # ---- the OTHER encoding as a full record (BASELINE.json configs[2] names the 2K frame with the HashGrid encoding;
        #      the headline is quoted on the reference's default, Frequency)
        oenc = 1 - enc
        O = Bench(torch, pkg, synth, oenc, nq, rank, world, args)
        n_o = max(10, min(args.steps, 50))
        o_ms, o_inf, o_tr, o_l = O.timed(n_o, 3)
        o_e2e_s, o_e2e = O.e2e(max(1, min(args.e2e_steps, n_o)))
        rec = {"name": f"BASELINE configs[2]: 2K frame, {ENC_NAME[oenc]}" if oenc == 1 else f"2K frame, {ENC_NAME[oenc]}",
               "config": config_dict(oenc, nq, 1, args.dp, args.query_order), "metric": METRIC, "unit": UNIT,
               "value": (nq + n_train) / (o_ms * 1e-3), "ms_per_step": o_ms, "ms_infer": o_inf, "ms_train_4_steps": o_tr, "steps": n_o,
               "queries_per_s": nq / (o_inf * 1e-3), "train_samples_per_s": n_train / (o_tr * 1e-3),
               "e2e": dict({"value": (nq + n_train) / o_e2e_s, "unit": UNIT}, **o_e2e), "gpu_launches": o_l,
               "rooflines": rooflines_for(oenc, nq, o_inf, o_tr, peaks, traffic, l2_peak)}
        O.close()
        del O
        torch.cuda.empty_cache()
        configs.append(rec)

    strong_4k = None
    if world > 1 and not strong and not args.no_configs:
        # ---- BASELINE.json configs[3]: ONE 4K frame (8 424 000 queries) batch-sharded over the ranks + training batches split
        lo, hi = distmod.shard_range(FRAME_4K_QUERIES, rank, world)
        S = Bench(torch, pkg, synth, enc, hi - lo, rank, world, args, width=3840, train_local=BATCH_SIZE // world, dp=dp_factory, strong=True)
        s_ms, s_inf, s_tr, _ = S.timed(max(10, min(args.steps, 50)), 3)
        strong_4k = {"workload": config_dict(enc, hi - lo, world, args.dp, args.query_order, FRAME_4K_QUERIES)["workload"],
                     "scaling": "strong", "ms_per_step": s_ms, "ms_infer": s_inf, "ms_train_4_steps": s_tr,
                     "value": (FRAME_4K_QUERIES + n_train) / (s_ms * 1e-3), "unit": UNIT, "queries_per_gpu": hi - lo}
        S.close()
        del S

    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
            "ms_per_step": ms_step, "higher_is_better": True, "scaling": "strong" if strong else "weak", "vs_baseline": None, "dtype": "f16",
            "data": "synthetic",
            "config": config_dict(enc, nq, world, args.dp, args.query_order, args.shard_frame if strong else 0),
            "queries_per_s": (args.shard_frame if strong else world * nq) / (t_inf * 1e-3),
            "train_samples_per_s": (n_train if strong else world * n_train) / max(t_tr * 1e-3, 1e-9),
            "ms_infer": t_inf, "ms_train_4_steps": t_tr,
            "frame_driver_ms": frame_ms, "clocks": clocks, "e2e": e2e, "gpu_launches": launches, "roofline": roofline,
            "rooflines": rl + extra_rl, "configs": configs}
    if strong_4k:
        line["strong_4k"] = strong_4k
    return line, rank, world, (q_np, tq_np, tt_np)


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
    line, rank, world, (q_np, tq_np, tt_np) = run_ours(args, enc)
    if rank == 0:
        if world == 1 and not args.no_cpu_baseline:
            orc = entry.load_oracle()
            synth = entry.load_package_synthetic()
            line["cpu_baseline"] = cpu_baseline_leg(orc, synth, enc, q_np, tq_np, tt_np, orc.ncpu())
        if world == 1 and not args.no_ref_gpu and args.queries == FRAME_QUERIES:
            # "beat the reference's kernels on the same box": its tcnn build for sm_100, both encodings, same frame shape
            line["reference_gpu"] = time_reference_gpu(enc, args.queries)
            ref_ms = line["reference_gpu"].get("ms_per_update_async")
            if ref_ms:
                line["reference_gpu"]["speedup_device_timed"] = ref_ms / line["ms_per_step"]
            for c in line["configs"]:
                oenc = 1 if "Hash" in c["config"]["workload"] else 0
                c["reference_gpu"] = time_reference_gpu(oenc, args.queries)
                r = c["reference_gpu"].get("ms_per_update_async")
                if r:
                    c["reference_gpu"]["speedup_device_timed"] = r / c["ms_per_step"]
        print(json.dumps(line), file=out, flush=True)
    if world > 1:
        import torch.distributed as dist
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
