#!/bin/bash
# Round-end measurement pass on one B200: parity suite, bench lines (both encodings), A/B builds, sweep, timeline, ncu.
T=${1:-f}
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/${T}_pytest.log 2>&1; echo "pytest exit $?"; tail -2 gpurun_out/${T}_pytest.log
timeout 400 python bench.py > gpurun_out/${T}_bench_frequency.json 2> gpurun_out/${T}_bench_frequency.err; echo "bench freq exit $?"
timeout 400 python bench.py --encoding hash > gpurun_out/${T}_bench_hash.json 2> gpurun_out/${T}_bench_hash.err; echo "bench hash exit $?"
timeout 600 python tools/ab_variants.py > gpurun_out/${T}_ab.log 2>&1; echo "ab exit $?"
timeout 400 python tools/sweep.py --out gpurun_out/${T}_sweep.json > gpurun_out/${T}_sweep.log 2>&1; echo "sweep exit $?"
timeout 200 python tools/train_timeline.py > gpurun_out/${T}_timeline.log 2>&1; cat gpurun_out/${T}_timeline.log
timeout 400 bash tools/profile_r01.sh frequency $T > gpurun_out/${T}_profile_f.log 2>&1; echo "profile f exit $?"
timeout 400 bash tools/profile_r01.sh hash $T > gpurun_out/${T}_profile_h.log 2>&1; echo "profile h exit $?"
cut -c1-400 gpurun_out/${T}_bench_frequency.json
