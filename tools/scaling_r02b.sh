#!/bin/bash
# 8-GPU evidence, trimmed to what the driver's own 1 -> 8 run does not produce (run with `gpurun --gpus 8 -- bash tools/scaling_r02b.sh`):
# the default line at N (weak + strong_4k + Hash record), the NCCL all-reduce variant, Hash as the headline, BASELINE configs[4] sweep.
N=${1:-8}; T=${2:-r02b}
mkdir -p gpurun_out
run() { P=$1; shift; python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $P "$@"; }
timeout 400 bash -c "$(declare -f run); N=$N; run 29611 bench.py --gpus $N --steps 50 --warmup 5 --no-ref-gpu" > gpurun_out/${T}_bench_$N.json 2> gpurun_out/${T}_bench_$N.err; echo "bench $N exit $?"
timeout 300 bash -c "$(declare -f run); N=$N; run 29623 bench.py --gpus $N --steps 50 --warmup 5 --dp nccl --no-configs --no-ref-gpu" > gpurun_out/${T}_bench_${N}_nccl.json 2> gpurun_out/${T}_bench_${N}_nccl.err; echo "bench nccl exit $?"
timeout 300 bash -c "$(declare -f run); N=$N; run 29635 bench.py --gpus $N --steps 50 --warmup 5 --encoding hash --no-configs --no-ref-gpu" > gpurun_out/${T}_bench_${N}_hash.json 2> gpurun_out/${T}_bench_${N}_hash.err; echo "bench hash exit $?"
timeout 500 bash -c "$(declare -f run); N=$N; run 29647 tools/sweep.py --out gpurun_out/${T}_sweep_$N.json --reps 8" > gpurun_out/${T}_sweep_$N.txt 2> gpurun_out/${T}_sweep_$N.err; echo "sweep exit $?"
python - <<PY
import json, glob
for f in sorted(glob.glob("gpurun_out/${T}_bench_*.json")):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
        print(f, d["n_gpus"], "ms/step %.4f" % d["ms_per_step"], "value %.3e" % d["value"], "infer %.4f train4 %.4f" % (d["ms_infer"], d["ms_train_4_steps"]),
              "e2e ms %.2f" % d["e2e"]["ms_per_step"], "pcie/rank", d["e2e"].get("pcie_gbs_per_rank"), "strong4k", (d.get("strong_4k") or {}).get("ms_per_step"))
    except Exception as e:
        print(f, "unreadable", e)
PY
tail -5 gpurun_out/${T}_sweep_$N.txt
