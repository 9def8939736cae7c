# quick A/B: tools/ab_variants.py over the default build + build/variants (no pytest), then the Frequency / Hash step timeline of each build
timeout 1200 python tools/ab_variants.py > gpurun_out/abq_variants.log 2>&1; echo "ab rc=$?"
python - <<'PY'
import json
for r in json.load(open("gpurun_out/ab_variants.json")):
    print(r["lib"], "infer0 %.1f rnd %.1f train0 %.1f | infer1 %.1f rnd %.1f train1 %.1f | exact %.4f %.4f" % (r["enc0_infer_us"], r["enc0_infer_random_2M_us"], r["enc0_train4_us"], r["enc1_infer_us"], r["enc1_infer_random_2M_us"], r["enc1_train4_us"], r["enc0_cornell_exact"], r["enc1_cornell_exact"]))
PY
for lib in "" build/variants/*.so; do echo "== ${lib:-default}"; NRC_B200_LIB=$lib timeout 300 python tools/train_timeline.py 2>&1 | grep -E "^enc|back-to-back"; done
