#!/bin/bash
# same-box A/B of two builds of the library inside the full step: tools/ab_lib.sh build/variants/a.so [build/variants/b.so]
A=$1; B=${2:-music_generation_b200/lib/libb200svs.so}
for lib in $A $B $A $B; do
  echo "== $lib"
  B200SVS_LIB=$lib python bench.py --steps 8 --warmup 3 --no-cpu-baseline --no-extras 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.readline())
st=d['stages_ms']
print(round(d['ms_per_step'],2), {k.split(':')[1] if ':' in k and 'untagged' not in k else k.split(':')[0]:round(v['ms'],1) for k,v in st.items() if v['ms']>1.4}, 'mhz', d['clocks']['sm_mhz'])"
done
