#!/bin/bash
# bench.py value/e2e for a list of decoder option strings (quick sweep, no CPU baseline)
for o in "$@"; do
  python bench.py --steps 10 --warmup 3 --skip-cpu-baseline --options ":fancy_upsampling=1 $o" 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('$o', 'ms/step %.3f'%d['ms_per_step'], 'img/s %.0f'%d['images_per_s'], 'e2e ms %.2f'%d['e2e']['ms_per_step'], 'device-out ms %.3f'%d['e2e_device_out']['ms_per_step'])"
done
