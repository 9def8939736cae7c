#!/bin/bash
# One GPU-box visit: parity tests, bench line, A/B of the encode builds, configs[4] sweep, ncu evidence (scratch orchestration).
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/a_pytest.log 2>&1; echo "pytest exit $?" | tee -a gpurun_out/a_pytest.log
tail -3 gpurun_out/a_pytest.log
timeout 400 python bench.py > gpurun_out/a_bench_frequency.json 2> gpurun_out/a_bench_frequency.err; echo "bench exit $?"
timeout 600 python tools/ab_variants.py > gpurun_out/a_ab.log 2>&1; echo "ab exit $?"; cat gpurun_out/a_ab.log | cut -c1-600
timeout 400 python tools/sweep.py --out gpurun_out/a_sweep.json > gpurun_out/a_sweep.log 2>&1; echo "sweep exit $?"; tail -50 gpurun_out/a_sweep.log
timeout 400 bash tools/profile_r01.sh frequency a > gpurun_out/a_profile.log 2>&1; echo "profile exit $?"
cut -c1-1500 gpurun_out/a_bench_frequency.json
