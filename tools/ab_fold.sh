#!/bin/bash
# A/B of time-pair folding at the 64-channel stage (CTA pairs make the folded N = 128 form's weights resident per CTA)
for kw in '{}' '{"generator_fold": {"32": 11, "64": [7, 11]}}' '{"generator_fold": {"32": 11, "64": 11}}' '{"generator_fold": {"32": 11, "64": [11]}}' '{}'; do
  echo "== $kw"
  python bench.py --steps 8 --warmup 3 --no-cpu-baseline --no-extras --model-kw "$kw" 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.readline())
st=d['stages_ms']
print(round(d['ms_per_step'],2), 'rb2', st['conv1d_tc:gen_rb2']['ms'], 'rb3', st['conv1d_tc:gen_rb3']['ms'], 'rb1', st['conv1d_tc:gen_rb1']['ms'], 'mhz', d['clocks']['sm_mhz'])"
done
