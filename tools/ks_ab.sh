#!/bin/bash
# tools/ks_ab.sh VAR v1 v2 ... -- key-switch-only bench line per value of an environment switch (run on the GPU box)
VAR=$1; shift
for v in "$@"; do
  env $VAR=$v python bench.py --ks-only --steps 6 --warmup 2 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read())
print('$VAR=$v', 'ks/s %.0f' % d['value'], 'ip_f64 us/launch %.1f' % d['roofline']['avg_launch_us'], 'frac %.3f' % d['roofline']['frac'], 'clk', d['clocks'])
print('   split', d.get('keyswitch_stage_split'))
"
done
