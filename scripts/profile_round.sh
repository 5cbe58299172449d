#!/bin/bash
# Per-round profile set (run on the GPU box AFTER tests and bench have passed without ncu):
#   launch lists (durations only) of the batch-64 / batch-1 decode step and of the DAC decoder, CUPTI timelines of the
#   step, and one `ncu --set full` capture of the dominant kernels.  usage: scripts/profile_round.sh <tag>
tag=${1:-rXX}
mkdir -p gpurun_out
for b in 64 1; do
  timeout 300 python scripts/profile_step.py --batch $b --ctx 1432 > gpurun_out/ps$b.log 2>&1 || { echo "profile_step $b failed"; exit 1; }
  timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv \
    --log-file gpurun_out/${tag}_launches_decode_b${b}_ctx1432.csv python scripts/profile_step.py --batch $b --ctx 1432 > gpurun_out/ncu$b.log 2>&1
  echo "launch list b$b rc=$?"
  timeout 300 python scripts/timeline_step.py --batch $b --layers 1 2>/dev/null | grep -v Warn > gpurun_out/${tag}_timeline_b${b}.txt
done
timeout 200 python scripts/profile_dac.py 8 860 > gpurun_out/pd.log 2>&1 || { echo "profile_dac failed"; exit 1; }
timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv \
  --log-file gpurun_out/${tag}_launches_dac_b8_t860.csv python scripts/profile_dac.py 8 860 > gpurun_out/ncu_dac.log 2>&1
echo "launch list dac rc=$?"
# full captures: attention + the four GEMMs of a layer at batch 64 (skip layer 0), two DAC layers (96-ch k7 and k1)
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"attn_decode|gemm_skinny|resid_ln" --profile-from-start off \
  -s 8 -c 7 -o gpurun_out/${tag}_full_decode_b64 -f python scripts/profile_step.py --batch 64 --ctx 1432 > gpurun_out/ncu_full64.log 2>&1
echo "full decode rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:dac_conv --profile-from-start off -s 23 -c 2 \
  -o gpurun_out/${tag}_full_dac96 -f python scripts/profile_dac.py 8 860 > gpurun_out/ncu_fulldac.log 2>&1
echo "full dac rc=$?"
ls -la gpurun_out/${tag}_*
