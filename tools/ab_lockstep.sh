#!/bin/bash
# A/B on one box: the lock-step and host-to-host legs with the two latest engine changes on and off.
for cfg in "" "RMEM_LN_COPY=0" "RMEM_LABEL_ALIAS=0" "RMEM_LN_COPY=0 RMEM_LABEL_ALIAS=0"; do
  env $cfg python bench.py --steps 60 --warmup 12 --no-cpu-baseline 2>/dev/null | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read())
print('$cfg', 'single', round(d['value']), 'e2e', round(d['e2e']['value']), 'u8', round(d['e2e_u8']['value']), 'lock', {k:round(v) for k,v in d['multi_sequence']['lockstep_by_S'].items()}, 'cfg4', round(d['configs']['4']['value']), 'cfg5', round(d['configs']['5']['value']))"
done
