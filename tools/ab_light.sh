#!/bin/bash
# tools/ab_light.sh NAME...  (GPU box): light_stress + a short bench per variant library
cd "$(dirname "$0")/.."
cp ash_renderer_b200/libashcuda.so /tmp/libashcuda_orig.so
for n in "$@"; do
  cp ash_renderer_b200/variants/libashcuda_$n.so ash_renderer_b200/libashcuda.so
  echo "== $n"; python tools/light_stress.py | tr -d "\n " | python -c "import json,sys; d=json.loads(sys.stdin.read()); print({k:(round(v['prepass_and_cells_us'],1), round(v['tiles_us'],1)) for k,v in d.items()})"
  python bench.py --steps 5 --warmup 3 --cpu-views 0 --precision fast --sustain-s 0 --total-views 0 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('  views/s %.0f' % d['value'])"
done
cp /tmp/libashcuda_orig.so ash_renderer_b200/libashcuda.so
