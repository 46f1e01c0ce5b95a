#!/bin/bash
# tools/run_profile.sh TAG  (on the GPU box): parity tests, a clean bench, then the ncu launch list and
# --set full captures of the top kernels of the SAME command.  Outputs under gpurun_out/.
TAG=${1:-x}
cd "$(dirname "$0")/.."
python -m pytest tests -m gpu -q 2>&1 | tail -4 > gpurun_out/pytest_gpu_$TAG.log; cat gpurun_out/pytest_gpu_$TAG.log
python bench.py > gpurun_out/bench_$TAG.json 2> gpurun_out/bench_$TAG.err; tail -c 400 gpurun_out/bench_$TAG.json
python bench.py --impl reference --steps 2 --warmup 0 > gpurun_out/bench_ref_$TAG.json 2> gpurun_out/bench_ref_$TAG.err
B="python bench.py --steps 2 --warmup 3 --streams 1 --views 2 --distinct 2 --cpu-views 0 --sustain-s 0 --total-views 0"
$B > gpurun_out/plain_$TAG.json 2> gpurun_out/plain_$TAG.err && \
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 700 --csv --log-file gpurun_out/launches_$TAG.csv $B > gpurun_out/ncu_$TAG.log 2>&1
# launch index captured per kernel: k_upsample_pair 5 = a fused upsample + tonemap instance (2 launches per view);
# k_downsample_box 6 = a 4K -> 1080p level (3 launches per view)
for KS in k_taa_tol:5 k_upsample_pair:5 k_downsample_box:6 k_depth_pyramid:5 k_light_prepass:5 k_light_cell_tiles:5 k_occlusion_cull:5; do
  K=${KS%%:*}; S=${KS##*:}
  ncu --set full --clock-control none --import-source on -k regex:$K -s $S -c 1 -f -o gpurun_out/full_${TAG}_$K $B > gpurun_out/ncu_full_${TAG}_$K.log 2>&1
done
ls -la gpurun_out | tail -20
