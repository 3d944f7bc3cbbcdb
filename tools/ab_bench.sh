#!/bin/bash
# A/B of SynthesizerTrn constructor knobs on one GPU box: tools/ab_bench.sh '{}' '{"generator_mrf": "fp32"}' ...
# Each setting is benched twice, interleaved (box-to-box and run-to-run noise is +-3 %).
mkdir -p gpurun_out
i=0
for rep in 1 2; do
  for kw in "$@"; do
    i=$((i+1))
    python bench.py --steps 3 --warmup 3 --no-cpu-baseline --model-kw "$kw" > gpurun_out/ab_$i.json 2> gpurun_out/ab_$i.err || tail -3 gpurun_out/ab_$i.err
    python - "$kw" gpurun_out/ab_$i.json <<'PY'
import json,sys
d=json.loads(open(sys.argv[2]).read().strip().splitlines()[-1])
st=d['stages_ms']; g=lambda k:(st[k]['ms'] if isinstance(st[k],dict) else st[k])
print(sys.argv[1], round(d['value']), round(d['ms_per_step'],1), d['clocks']['sm_mhz'], 'rb0-3', [round(g(f'conv1d_tc:gen_rb{i}'),1) for i in range(4)], 'specdec', round(g('conv1d_tc:specdec'),2))
PY
  done
done
