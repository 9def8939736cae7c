#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_train_gpu.py tests/test_frame.py tests/test_golden.py -m gpu -x -q 2>&1 | tail -5
timeout 300 python tools/train_timeline.py 2>&1 | tee gpurun_out/e_timeline.log
timeout 600 python tools/ab_variants.py 2>&1 | cut -c1-100,380-480 | tee gpurun_out/e_ab.log
