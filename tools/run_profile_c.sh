set -x
python -m pytest tests -m gpu -q 2>&1 | tail -8 > gpurun_out/pytest_gpu.log; cat gpurun_out/pytest_gpu.log
B="python bench.py --steps 2 --warmup 3 --streams 1 --views 2 --distinct 2 --cpu-views 0"
$B > gpurun_out/plain_c.json 2> gpurun_out/plain_c.err && \
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_c.csv $B > gpurun_out/ncu_c.log 2>&1
for K in k_taa_fast k_upsample_fast k_light_cells k_downsample_fast; do
  ncu --set full --clock-control none --import-source on -k regex:$K -s 3 -c 1 -f -o gpurun_out/full_c_$K $B > gpurun_out/ncu_full_$K.log 2>&1
done
ls -la gpurun_out
