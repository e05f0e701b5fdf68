#!/bin/bash
# on the GPU box: pose tests once (main build), then the motion-only BA bench for every build_exp/libb2ba_*.so
timeout 300 python -m pytest tests/test_pose.py -x -q -m gpu 2>&1 | tail -3
cp visualodometry_b200/libb2ba.so /tmp/libb2ba_keep.so
for f in build_exp/libb2ba_*.so; do
  cp $f visualodometry_b200/libb2ba.so
  echo -n "$f: "
  timeout 200 python bench.py --workload pnp 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['ms_per_step'], round(d['roofline']['frac'],4), d['e2e']['ms_per_step'])"
done
cp /tmp/libb2ba_keep.so visualodometry_b200/libb2ba.so
