#!/bin/bash
# front-end cost experiments (GPU box)
cd /root/repo
run() { ASH_BENCH_STAGES=$1 python bench.py --steps 5 --warmup 3 --cpu-views 0 --precision fast --sustain-s 0 --total-views 0 --streams 8 --graph-replay ${2:-1} 2>gpurun_out/st.err | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('  views/s %.0f  us/view %.2f  launches/view %.1f' % (d['value'], 1e6/d['value'], d['gpu_launches']/160))" || tail -3 gpurun_out/st.err; }
cp ash_renderer_b200/libashcuda.so /tmp/orig.so
for m in 0x01 0x03 0x05 0x07; do echo "== base stages $m"; run $m; done
cp ash_renderer_b200/variants/libashcuda_skipsmall.so ash_renderer_b200/libashcuda.so
echo "== skipsmall stages 0x07 (graph replay off: the store node is absent)"; run 0x07 0
cp /tmp/orig.so ash_renderer_b200/libashcuda.so
echo "== base stages 0x07 graph replay off"; run 0x07 0
