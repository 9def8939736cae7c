#!/bin/bash
mkdir -p gpurun_out
for lib in build/variants/*.so; do
  NRC_B200_LIB=$lib timeout 300 python -m pytest tests/test_infer_gpu.py tests/test_frame.py -m gpu -x -q 2>&1 | tail -1 | sed "s#^#$lib: #"
done
timeout 600 python tools/ab_variants.py 2>&1 | cut -c1-700 | tee gpurun_out/d_ab.log
timeout 400 bash tools/profile_r01.sh frequency d > gpurun_out/d_profile.log 2>&1; echo "profile exit $?"
