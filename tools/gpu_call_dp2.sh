# 2-GPU check of the data-parallel paths: parity tests (fused == NCCL == single GPU, one process driving two devices) + short benches
run() { N=$1; shift; python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29500 + RANDOM % 200)) "$@"; }
timeout 600 python -m pytest tests/test_dp_gpu.py tests/test_multi_device.py -m gpu -q --no-header -rfs > gpurun_out/dp2_pytest.log 2>&1; echo "dp pytest rc=$?"; tail -4 gpurun_out/dp2_pytest.log
timeout 400 bash -c "$(declare -f run); run 2 bench.py --gpus 2 --steps 50 --warmup 5 --no-configs --no-ref-gpu" > gpurun_out/dp2_bench_freq.json 2> gpurun_out/dp2_bench_freq.err; echo "bench freq rc=$?"
timeout 400 bash -c "$(declare -f run); run 2 bench.py --gpus 2 --steps 50 --warmup 5 --no-configs --no-ref-gpu --encoding hash" > gpurun_out/dp2_bench_hash.json 2> gpurun_out/dp2_bench_hash.err; echo "bench hash rc=$?"
python - <<'PY'
import json
for f in ("gpurun_out/dp2_bench_freq.json", "gpurun_out/dp2_bench_hash.json"):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
        print(f, d["n_gpus"], "ms/step %.4f infer %.4f train4 %.4f e2e %.2f" % (d["ms_per_step"], d["ms_infer"], d["ms_train_4_steps"], d["e2e"]["ms_per_step"]), d.get("strong_4k"))
    except Exception as e:
        print(f, "unreadable", e); print(open(f.replace(".json", ".err")).read()[-1500:])
PY
