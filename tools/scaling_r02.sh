#!/bin/bash
# Multi-GPU evidence of round 2 on one box (run with `gpurun --gpus 8 -- bash tools/scaling_r02.sh r02 "2 4 8"`):
#   2-GPU parity tests (data-parallel paths, one process driving two devices), the 1 -> N scaling bench (weak + strong_4k extra),
#   the NCCL all-reduce variant and the Hash encoding at the largest N, BASELINE.json configs[4] sweep at the largest N.
T=${1:-r02}; NS=${2:-"2 4 8"}
mkdir -p gpurun_out
run() { N=$1; shift; python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29500 + N + RANDOM % 200)) "$@"; }
timeout 600 python -m pytest tests/test_dp_gpu.py tests/test_multi_device.py -m gpu -q --no-header -rfs -s > gpurun_out/${T}_dp_pytest.log 2>&1
echo "dp pytest exit $?"; tail -n 6 gpurun_out/${T}_dp_pytest.log
timeout 300 python bench.py --gpus 1 --steps 50 --warmup 5 --no-cpu-baseline --no-ref-gpu --no-configs > gpurun_out/${T}_bench_1.json 2> gpurun_out/${T}_bench_1.err; echo "bench 1 exit $?"
LAST=1
for N in $NS; do
  timeout 400 bash -c "$(declare -f run); run $N bench.py --gpus $N --steps 50 --warmup 5" > gpurun_out/${T}_bench_$N.json 2> gpurun_out/${T}_bench_$N.err; echo "bench $N exit $?"
  LAST=$N
done
timeout 400 bash -c "$(declare -f run); run $LAST bench.py --gpus $LAST --steps 50 --warmup 5 --dp nccl --no-configs" > gpurun_out/${T}_bench_${LAST}_nccl.json 2> gpurun_out/${T}_bench_${LAST}_nccl.err; echo "bench nccl exit $?"
timeout 400 bash -c "$(declare -f run); run $LAST bench.py --gpus $LAST --steps 50 --warmup 5 --encoding hash --no-configs" > gpurun_out/${T}_bench_${LAST}_hash.json 2> gpurun_out/${T}_bench_${LAST}_hash.err; echo "bench hash exit $?"
timeout 600 bash -c "$(declare -f run); run $LAST tools/sweep.py --out gpurun_out/${T}_sweep_${LAST}.json --reps 10" > gpurun_out/${T}_sweep_${LAST}.txt 2> gpurun_out/${T}_sweep_${LAST}.err; echo "sweep exit $?"
python - <<PY
import json, glob
for f in sorted(glob.glob("gpurun_out/${T}_bench_*.json")):
    try:
        d = json.load(open(f))
        print(f, d["n_gpus"], "ms/step %.4f" % d["ms_per_step"], "value %.3e" % d["value"], "infer %.4f train4 %.4f" % (d["ms_infer"], d["ms_train_4_steps"]),
              "e2e ms %.2f" % d["e2e"]["ms_per_step"], d.get("strong_4k", {}).get("ms_per_step"))
    except Exception as e:
        print(f, "unreadable", e)
PY
