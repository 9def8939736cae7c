#!/bin/bash
# ncu evidence for one encoding (round 2): launch list of the bench command + full captures of the query kernel and the
# training step kernel.  usage (on the GPU box, via gpurun): bash tools/profile_r02.sh frequency|hash <tag>
ENC=${1:-frequency}; TAG=${2:-r02}
CMD="python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-configs --no-ref-gpu --e2e-steps 1 --encoding $ENC"
$CMD > gpurun_out/${TAG}_plain_${ENC}.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/${TAG}_plain_${ENC}.log; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${TAG}_launches_${ENC}.csv $CMD > /dev/null 2>&1
ncu --set full --clock-control none --import-source on -k regex:nrc_infer -s 3 -c 1 -f -o gpurun_out/${TAG}_infer_${ENC} $CMD > /dev/null 2>&1
ncu --set full --clock-control none --import-source on -k regex:nrc_train_step -s 3 -c 1 -f -o gpurun_out/${TAG}_train_${ENC} $CMD > /dev/null 2>&1
ls -la gpurun_out/ | grep ${TAG}
