#!/bin/bash
# A/B timing of walk-kernel variants on one box: tools/ab_walk.sh var1 var2 ... (libraries shenqi_b200/csrc/variants/var_<name>.so)
cp shenqi_b200/libshenqi_b200.so /tmp/lib_keep.so
for v in "$@"; do
  cp shenqi_b200/csrc/variants/var_$v.so shenqi_b200/libshenqi_b200.so
  for rep in 1 2; do
    echo -n "$v " ; python tools/prof_step.py --ngrid 256 --steps 3 --no-pm | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['walk_ms'], d['interactions'])"
  done
done
cp /tmp/lib_keep.so shenqi_b200/libshenqi_b200.so
