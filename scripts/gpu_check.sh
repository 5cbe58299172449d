# GPU-box check used during development: kernel + model parity tests, then short benches (400 frames) under a few
# planner settings.  Usage: bash scripts/gpu_check.sh [env-setting ...]   e.g.  "ZB200_SK_MAXSPLIT=2"
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_kernels.py -x -q -m gpu -k "gemm" > gpurun_out/t_kernels.log 2>&1; echo "kernels rc=$?"; tail -4 gpurun_out/t_kernels.log
timeout 900 python -m pytest tests -x -q -m gpu --deselect tests/test_gpu_kernels.py > gpurun_out/t_rest.log 2>&1; echo "rest rc=$?"; tail -4 gpurun_out/t_rest.log
i=0
for setting in "" "$@"; do
  i=$((i+1))
  env $setting timeout 400 python bench.py --max-new-tokens 400 --steps 1 --warmup 1 --no-cpu-baseline --no-dac > gpurun_out/b_$i.log 2>&1
  echo "bench[$setting] rc=$?"
  tail -1 gpurun_out/b_$i.log | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['value'], d['decode_step']['ms_per_frame'], d['latency_batch1']['p50_ms'])"
done
