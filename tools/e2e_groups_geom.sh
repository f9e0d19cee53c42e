#!/bin/bash
# host-pointer call (emb_lookup_ex, C2): group-size growth ratio (percent) x group count
for r in 100 115 130 150; do for g in 5 7; do
  echo "EMB_GROUP_RATIO=$r EMB_GROUPS=$g"
  EMB_GROUP_RATIO=$r EMB_GROUPS=$g timeout 300 python bench.py --steps 20 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); e=d['e2e']; print('   e2e ms/step', round(e['ms_per_step'],3), 'G lookups/s', round(e['value']/1e9,2), e['last_call_ms'])"
done; done
