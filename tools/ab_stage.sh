#!/bin/bash
# A/B of library variants on one box, all stage timers: tools/ab_stage.sh var1 var2 ... (shenqi_b200/csrc/variants/var_<name>.so)
cp shenqi_b200/libshenqi_b200.so /tmp/lib_keep.so
for v in "$@"; do
  cp shenqi_b200/csrc/variants/var_$v.so shenqi_b200/libshenqi_b200.so
  for rep in 1 2; do
    echo -n "$v " ; python tools/prof_step.py --ngrid 256 --steps 3 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print({k: round(v, 3) for k, v in d['stages'].items() if k in ('sort_ms', 'pm_paint_ms', 'pm_readout_ms', 'pm_fft_ms', 'walk_ms')})"
  done
done
cp /tmp/lib_keep.so shenqi_b200/libshenqi_b200.so
