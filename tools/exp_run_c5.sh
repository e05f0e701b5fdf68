#!/bin/bash
# on the GPU box: C5 strong-scaling record (ms per LM iteration + phases) for every build_exp/libb2ba_*.so
cp visualodometry_b200/libb2ba.so /tmp/libb2ba_keep.so
for f in build_exp/libb2ba_*.so; do
  cp $f visualodometry_b200/libb2ba.so
  echo -n "$f: "
  timeout 300 python tools/c5_scaling.py ${C5_MODES:-random local} 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
for m in d:
    if m=='n_gpus': continue
    x=d[m]; print(m, round(x['ms_per_step'],2), x['pcg_iterations'], {k:round(v,3) for k,v in x['phases_ms_per_lm_iteration'].items()}, end=' | ')
print()"
done
cp /tmp/libb2ba_keep.so visualodometry_b200/libb2ba.so
