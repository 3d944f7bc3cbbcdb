#!/bin/bash
# On the GPU box: bench every build/variants/*.so (named on the command line) back to back on the same GPU.
mkdir -p gpurun_out
cp music_generation_b200/lib/libb200svs.so /tmp/lib_base.so
for v in base "$@" base; do
  if [ "$v" = base ]; then cp /tmp/lib_base.so music_generation_b200/lib/libb200svs.so; else cp build/variants/$v.so music_generation_b200/lib/libb200svs.so; fi
  python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/var_$v.json 2> gpurun_out/var_$v.err || tail -3 gpurun_out/var_$v.err
  python - "$v" <<'PY'
import json,sys
v=sys.argv[1]
d=json.loads(open(f'gpurun_out/var_{v}.json').read().strip().splitlines()[-1])
st=d['stages_ms']; g=lambda k:(st[k]['ms'] if isinstance(st[k],dict) else st[k])
print(v, round(d['value']), round(d['ms_per_step'],1), d['clocks']['sm_mhz'], 'rb0-3', [round(g(f'conv1d_tc:gen_rb{i}'),1) for i in range(4)], 'attn', round(g('attention_tc:untagged'),1), 'flow', round(g('conv1d_tc:flow'),2), 'specdec', round(g('conv1d_tc:specdec'),2), 'f0', round(g('conv1d_tc:f0'),2))
PY
done
cp /tmp/lib_base.so music_generation_b200/lib/libb200svs.so
