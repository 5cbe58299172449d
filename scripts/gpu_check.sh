mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_kernels.py -x -q -m gpu -k "gemm" > gpurun_out/t_kernels.log 2>&1; echo "kernels rc=$?"; tail -15 gpurun_out/t_kernels.log
timeout 900 python -m pytest tests -x -q -m gpu --deselect tests/test_gpu_kernels.py > gpurun_out/t_rest.log 2>&1; echo "rest rc=$?"; tail -15 gpurun_out/t_rest.log
timeout 400 python bench.py --max-new-tokens 400 --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/b_new.log 2>&1; echo "bench rc=$?"; tail -1 gpurun_out/b_new.log | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['value'], d['decode_step'], d['latency_batch1'])"
