#!/bin/bash
# DAC decoder: parity tests, end-to-end timing, then the per-layer launch list (ncu, durations only).
# usage: scripts/dac_layers.sh <tag>
tag=${1:-x}
timeout 300 python -m pytest tests/test_gpu_dac.py -x -q -m gpu 2>&1 | tail -4
timeout 120 python scripts/time_dac.py 8 860 || exit 1
timeout 120 python scripts/time_dac.py 64 2580 || exit 1
timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv \
  --log-file gpurun_out/launches_dac_$tag.csv python scripts/profile_dac.py 8 860 > gpurun_out/ncu_dac_$tag.log 2>&1
python - <<PY
import csv
rows=list(csv.reader(open('gpurun_out/launches_dac_$tag.csv')))
for i,r in enumerate(rows):
    if 'Kernel Name' in r: h=r; start=i; break
mi=h.index('Metric Value')
v=[float(r[mi].replace(',',''))/1e6 for r in rows[start+2:]]
print(' '.join('%.3f'%x for x in v)); print('total %.2f ms'%sum(v))
PY
