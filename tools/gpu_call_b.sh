#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q > gpurun_out/b_pytest_full.log 2>&1; echo "full pytest exit $?"; tail -8 gpurun_out/b_pytest_full.log
for i in 1 2 3 4; do
  timeout 200 python -m pytest tests/test_golden.py -m gpu -q -k "loss_curve and hash" -s 2>&1 | grep -E "AssertionError: \(|passed|failed|\[hash\]" | cut -c1-300 | sed "s/^/v3 run $i: /"
  NRC_B200_LIB=build/variants/libnrc_b200_v0.so timeout 200 python -m pytest tests/test_golden.py -m gpu -q -k "loss_curve and hash" -s 2>&1 | grep -E "AssertionError: \(|passed|failed|\[hash\]" | cut -c1-300 | sed "s/^/v0 run $i: /"
done 2>&1 | tee gpurun_out/b_hash_flaky.log
