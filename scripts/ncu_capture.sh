mkdir -p gpurun_out
timeout 300 python scripts/profile_step.py --batch 64 --ctx 1432 > gpurun_out/ps64.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:attn_decode --profile-from-start off -c 2 -o gpurun_out/prof_attn_b64 -f python scripts/profile_step.py --batch 64 --ctx 1432 > gpurun_out/ncu_attn.log 2>&1; echo "ncu attn rc=$?"
ncu --set full --clock-control none --import-source on -k regex:gemm_skinny --profile-from-start off -s 2 -c 3 -o gpurun_out/prof_gemm_b64 -f python scripts/profile_step.py --batch 64 --ctx 1432 > gpurun_out/ncu_gemm.log 2>&1; echo "ncu gemm rc=$?"
ls -la gpurun_out/*.ncu-rep
