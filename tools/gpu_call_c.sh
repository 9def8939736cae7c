#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_infer_gpu.py tests/test_frame.py -m gpu -x -q 2>&1 | tail -5
timeout 600 python tools/ab_variants.py 2>&1 | cut -c1-700 | tee gpurun_out/c_ab.log
timeout 900 python -m pytest tests -m gpu -q 2>&1 | tail -5
