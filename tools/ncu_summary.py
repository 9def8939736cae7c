"""Summarises ncu artefacts brought back in gpurun_out/ into small text files for profiles/.
usage: python tools/ncu_summary.py launches <csv>            -> per-kernel totals and shares
       python tools/ncu_summary.py rep <file.ncu-rep>        -> key counters + SASS opcode mix
       python tools/ncu_summary.py traffic <file.ncu-rep> <key> <summary.txt> [json]
                                                             -> dram__bytes_read.sum + dram__bytes_write.sum (and, key + "_l2",
                                                                lts__t_sectors.sum * 32) of the captured launch into
                                                                profiles/r02_traffic.json, which bench.py reads for roofline.traffic"""
import collections
import csv
import io
import subprocess
import sys

KEYS = ["gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
        "sm__cycles_elapsed.max", "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__pipe_tensor_cycles_active_realtime.avg.pct_of_peak_sustained_elapsed",
        "sm__pipe_tensor_subpipe_hmma_cycles_active_realtime.avg", "sm__inst_executed_pipe_tmem.avg.pct_of_peak_sustained_active",
        "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_bytes.sum", "lts__t_sector_hit_rate.pct",
        "lts__t_bytes.sum.per_second", "lts__t_sectors.sum", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum",
        "l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum", "l1tex__t_sectors_pipe_lsu_mem_global_op_ld_lookup_hit.sum",
        "l1tex__t_sector_hit_rate.pct", "l1tex__data_pipe_lsu_wavefronts.sum",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
        "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio"]


def launches(path):
    rows = [r for r in csv.reader(open(path)) if len(r) > 10]
    h = rows[0]
    ki, vi = h.index("Kernel Name"), h.index("Metric Value")
    agg = collections.defaultdict(list)
    for r in rows[1:]:
        agg[r[ki].split("(")[0][-60:]].append(float(r[vi].replace(",", "")))
    tot = sum(sum(v) for v in agg.values())
    print(f"# {path}: gpu__time_duration.sum per launch (ncu, cold cache, serialised: compare SHARES)")
    for k, v in sorted(agg.items(), key=lambda kv: -sum(kv[1])):
        print(f"{k:62s} n={len(v):4d} mean={sum(v) / len(v) / 1000:9.2f} us  max={max(v) / 1000:9.2f} us  "
              f"total={sum(v) / 1e6:8.3f} ms  share={100 * sum(v) / tot:5.1f}%")


def rep(path):
    raw = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    h = rows[0]
    for r in rows[2:]:
        name = r[h.index("Kernel Name")] if "Kernel Name" in h else "?"
        print(f"# {path}\n# kernel: {name}")
        for k in KEYS:
            # some counters carry a section prefix in the raw page ("TPC.TriageCompute.sm__pipe_tensor_..."): match the tail
            idx = [i for i, name in enumerate(h) if name == k or name.endswith("." + k)]
            if idx:
                i = idx[0]
                print(f"{k:90s} {r[i]:>18s} {rows[1][i]}")
    src = subprocess.run(["ncu", "-i", path, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(src)))
    hi = next(i for i, r in enumerate(rows) if "Source" in r and "Instructions Executed" in r)
    h = rows[hi]
    ai, ii, si = h.index("Source"), h.index("Instructions Executed"), h.index("# Samples")
    ops, samp, tot = collections.Counter(), collections.Counter(), 0
    for r in rows[hi + 1:]:
        if len(r) <= ii or not r[ai]:
            continue
        toks = r[ai].split()
        op = (toks[1] if toks[0].startswith("@") and len(toks) > 1 else toks[0]).split(".")[0]
        n = int(r[ii] or 0)
        ops[op] += n
        samp[op] += int(r[si] or 0)
        tot += n
    print(f"# SASS opcode mix (warp-level instructions executed, total {tot}); samples = warp-stall samples on that opcode")
    for k, v in ops.most_common(24):
        print(f"{k:12s} {v:12d} {100 * v / tot:5.1f}%  samples {samp[k]}")


def _to_bytes(value, unit):
    v = float(value.replace(",", ""))
    u = unit.strip().lower()
    return v * {"byte": 1, "kbyte": 1e3, "mbyte": 1e6, "gbyte": 1e9, "sector": 32}.get(u, 1)


def traffic(path, key, source, out=None):
    import json
    import os
    out = out or os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "profiles", "r02_traffic.json")
    raw = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    h, units, r = rows[0], rows[1], rows[2]

    def get(name):
        i = [k for k, n in enumerate(h) if n == name or n.endswith("." + name)][0]
        return _to_bytes(r[i], units[i])
    d = json.load(open(out)) if os.path.exists(out) else {}
    d[key] = {"bytes": get("dram__bytes_read.sum") + get("dram__bytes_write.sum"), "source": source,
              "what": "dram__bytes_read.sum + dram__bytes_write.sum of one launch (ncu --set full)"}
    d[key + "_l2"] = {"bytes": get("lts__t_sectors.sum"), "source": source, "what": "lts__t_sectors.sum x 32 B of one launch (ncu --set full)"}
    json.dump(d, open(out, "w"), indent=1)
    print(key, d[key], d[key + "_l2"])


if __name__ == "__main__":
    {"launches": launches, "rep": rep, "traffic": traffic}[sys.argv[1]](*sys.argv[2:])
