# A/B call: GPU parity tests on the default build (the no-alloc test also alone in a fresh process), tools/ab_variants.py over
# build/variants, the phase timeline of the default and of the HEAD build
timeout 900 python -m pytest tests -m gpu -x -q --no-header -rfs > gpurun_out/ab_pytest.log 2>&1; echo "pytest rc=$?"; tail -6 gpurun_out/ab_pytest.log
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q --no-header -rfs > gpurun_out/ab_pytest_noalloc_alone.log 2>&1; echo "noalloc alone rc=$?"; tail -4 gpurun_out/ab_pytest_noalloc_alone.log
timeout 900 python tools/ab_variants.py > gpurun_out/ab_variants.log 2>&1; echo "ab rc=$?"
python - <<'PY'
import json
for r in json.load(open("gpurun_out/ab_variants.json")):
    print(r["lib"], "infer0 %.1f train0 %.1f | infer1 %.1f rnd %.1f train1 %.1f | exact %.4f %.4f" % (r["enc0_infer_us"], r["enc0_train4_us"], r["enc1_infer_us"], r["enc1_infer_random_2M_us"], r["enc1_train4_us"], r["enc0_cornell_exact"], r["enc1_cornell_exact"]))
PY
timeout 300 python tools/train_timeline.py > gpurun_out/ab_timeline.txt 2>&1; cat gpurun_out/ab_timeline.txt
NRC_B200_LIB=build/variants/libnrc_b200_head.so timeout 300 python tools/train_timeline.py > gpurun_out/ab_timeline_head.txt 2>&1; cat gpurun_out/ab_timeline_head.txt
