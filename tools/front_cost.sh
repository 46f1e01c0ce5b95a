#!/bin/bash
# tools/front_cost.sh (GPU box): what each group of stages costs per view WITH other views' kernels beside it (8 streams, 32 views
# per step): bench.py restricted to a stage mask through ASH_BENCH_STAGES.  0x01 Hi-Z, 0x02 cull, 0x04 lights, 0x38 post chain.
cd "$(dirname "$0")/.."
run() { ASH_BENCH_STAGES=$1 python bench.py --steps 5 --warmup 3 --cpu-views 0 --precision fast --sustain-s 0 --total-views 0 --streams 8 2>gpurun_out/st.err | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('  views/s %.0f  us/view %.2f  launches/view %.1f' % (d['value'], 1e6/d['value'], d['gpu_launches']/160))" || tail -3 gpurun_out/st.err; }
for m in ${@:-0x01 0x03 0x05 0x07 0x38 0x3F}; do echo "== stages $m"; run $m; done
