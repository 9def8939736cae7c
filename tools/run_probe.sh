#!/bin/bash
# Runs every probe case in its own process (a faulting case must not poison the rest).
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
{
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv
if [ "$1" != "rate" ]; then
for c in ss_k ss_k80 ss_n16 ts ts_k80 ts_n16 mn_m128 mn_m64 mn_m64_n80; do
  timeout 60 build/umma_probe func $c; echo "exit=$?"
done
fi
timeout 120 build/umma_probe rate; echo "exit=$?"
} > gpurun_out/probe.log 2>&1
tail -60 gpurun_out/probe.log
