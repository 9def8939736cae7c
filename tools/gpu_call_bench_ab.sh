# bench-level A/B (the interplay of the query kernel and the training launch inside one update): default build vs build/variants/*.so
timeout 300 python -m pytest tests/test_infer_gpu.py tests/test_frame.py -m gpu -x -q --no-header 2>&1 | tail -2
for lib in "" build/variants/*.so; do for enc in frequency hash; do
  NRC_B200_LIB=$lib timeout 300 python bench.py --steps 100 --warmup 10 --encoding $enc --no-configs --no-cpu-baseline --no-ref-gpu --e2e-steps 2 2>/dev/null | python -c "
import sys, json
d = json.loads(sys.stdin.read().strip().splitlines()[-1]); print('${lib:-default}', '$enc', 'ms/step %.4f infer %.4f train4 %.4f frame_driver %.4f' % (d['ms_per_step'], d['ms_infer'], d['ms_train_4_steps'], d.get('frame_driver_ms') or 0))"
done; done
