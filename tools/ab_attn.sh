#!/bin/bash
# A/B of attention_ts (Q / P operands in TMEM) against attention_tc inside the full step
for m in 0 2 0 2; do
  echo "== B200SVS_ATTN_TS=$m"
  B200SVS_ATTN_TS=$m python bench.py --steps 8 --warmup 3 --no-cpu-baseline --no-extras 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.readline())
st=d['stages_ms']
print(round(d['ms_per_step'],2), {k:v['ms'] for k,v in st.items() if 'attention' in k or k.endswith(':f0') or k.endswith(':mel')}, 'mhz', d['clocks']['sm_mhz'])"
done
