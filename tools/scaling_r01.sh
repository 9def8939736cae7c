#!/bin/bash
# Multi-GPU bench lines for profiles/r01_scaling/: usage (via gpurun --gpus N): bash tools/scaling_r01.sh N
N=${1:-8}
run() { python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29500 + RANDOM % 400)) bench.py --gpus $N --steps 100 --warmup 5 --no-cpu-baseline "$@"; }
mkdir -p gpurun_out/scaling
run                      > gpurun_out/scaling/dp${N}_fused.json        2> gpurun_out/scaling/dp${N}_fused.err
run --shard-frame 8424000 > gpurun_out/scaling/strong4k_${N}.json       2> gpurun_out/scaling/strong4k_${N}.err
if [ "$N" = "8" ]; then
run --encoding hash      > gpurun_out/scaling/dp${N}_hash.json         2> gpurun_out/scaling/dp${N}_hash.err
true
fi
for f in gpurun_out/scaling/*_${N}*.json gpurun_out/scaling/dp${N}_*.json; do python - "$f" <<'PY'
import json, sys
try:
    d = json.load(open(sys.argv[1]))
    print(sys.argv[1], "value %.4g" % d["value"], "ms_per_step %.4f" % d["ms_per_step"], "infer %.4f" % d["ms_infer"], "train %.4f" % d["ms_train_4_steps"], d["scaling"], d["config"]["parallelism"])
except Exception as e:
    print(sys.argv[1], "FAILED", e)
PY
done
