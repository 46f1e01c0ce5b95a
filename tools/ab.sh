#!/bin/bash
# tools/ab.sh NAME...  (on the GPU box) run bench.py once per variant library; summaries to gpurun_out/ab_NAME.json
cd "$(dirname "$0")/.."
cp ash_renderer_b200/libashcuda.so /tmp/libashcuda_orig.so
for n in "$@"; do
  cp ash_renderer_b200/variants/libashcuda_$n.so ash_renderer_b200/libashcuda.so
  python bench.py --steps 5 --warmup 3 --cpu-views 0 --precision fast > gpurun_out/ab_$n.json 2> gpurun_out/ab_$n.err || tail -3 gpurun_out/ab_$n.err
  echo "== $n"; python tools/show_bench.py gpurun_out/ab_$n.json
done
cp /tmp/libashcuda_orig.so ash_renderer_b200/libashcuda.so
