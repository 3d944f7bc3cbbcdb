#!/bin/bash
# A/B of library environment knobs on one GPU box: tools/ab_env.sh "X=1" "B200SVS_TC_LEAN=0" ...
# Each setting is benched twice, interleaved (box-to-box and run-to-run noise is +-3 %).
mkdir -p gpurun_out
i=0
for rep in 1 2; do
  for kv in "$@"; do
    i=$((i+1))
    env $kv python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/abe_$i.json 2> gpurun_out/abe_$i.err || tail -3 gpurun_out/abe_$i.err
    python - "$kv" gpurun_out/abe_$i.json <<'PY'
import json,sys
d=json.loads(open(sys.argv[2]).read().strip().splitlines()[-1])
st=d['stages_ms']; g=lambda k:(st[k]['ms'] if isinstance(st[k],dict) else st[k])
print(sys.argv[1], round(d['value']), round(d['ms_per_step'],1), d['clocks']['sm_mhz'], 'rb0-3', [round(g(f'conv1d_tc:gen_rb{i}'),1) for i in range(4)])
PY
  done
done
