#!/bin/bash
# A/B of the fused ResBlock-pair kernel at the 32-channel stage inside the full step
for kw in '{"generator_pair_fused": false}' '{}' '{"generator_pair_fused": false}' '{}'; do
  echo "== $kw"
  python bench.py --steps 8 --warmup 3 --no-cpu-baseline --no-extras --model-kw "$kw" 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.readline())
st=d['stages_ms']
print(round(d['ms_per_step'],2), 'rb3', {k:v['ms'] for k,v in st.items() if 'rb3' in k or 'rb2' in k}, 'mhz', d['clocks']['sm_mhz'])"
done
