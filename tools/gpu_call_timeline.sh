# phase timeline of the default build and of every build/variants/*.so (timing experiments)
for lib in "" build/variants/*.so; do echo "== ${lib:-default}"; NRC_B200_LIB=$lib timeout 300 python tools/train_timeline.py 2>&1 | grep -E "^enc|back-to-back"; done
