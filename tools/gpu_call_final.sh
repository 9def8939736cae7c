# final 1-GPU measurement pass: GPU tests, the driver's bench command (reference GPU build included), the reference arm, ncu evidence
timeout 900 python -m pytest tests -m gpu -x -q --no-header -rfs > gpurun_out/final_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/final_pytest.log
timeout 900 python bench.py > gpurun_out/final_bench.json 2> gpurun_out/final_bench.err; echo "bench rc=$?"
timeout 900 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/final_bench_reference.json 2> gpurun_out/final_bench_reference.err; echo "reference arm rc=$?"
bash tools/profile_r02.sh frequency r02b
bash tools/profile_r02.sh hash r02b
