# Hash: four steps in one cooperative launch (default) vs one launch per batch (NRC_B200_HASH_MULTI=0), same box
timeout 900 python -m pytest tests -m gpu -x -q --no-header -rfs > gpurun_out/hm_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/hm_pytest.log
for m in 1 0 1 0; do
  NRC_B200_HASH_MULTI=$m NRC_AB_CHILD=1 timeout 600 python tools/ab_variants.py 2>&1 | grep "^AB " | python -c "
import sys, json
r = json.loads(sys.stdin.read()[3:]); print('HASH_MULTI=$m', 'train0 %.1f train1 %.1f infer1 %.1f' % (r['enc0_train4_us'], r['enc1_train4_us'], r['enc1_infer_us']))"
done
timeout 600 python bench.py --steps 50 --warmup 5 --encoding hash --no-configs --no-ref-gpu --no-cpu-baseline > gpurun_out/hm_bench_hash.json 2> gpurun_out/hm_bench_hash.err; echo "bench rc=$?"
python -c "
import json; d = json.loads(open('gpurun_out/hm_bench_hash.json').read().strip().splitlines()[-1]); print('hash bench ms/step %.4f infer %.4f train4 %.4f launches %d' % (d['ms_per_step'], d['ms_infer'], d['ms_train_4_steps'], d['gpu_launches']))"
