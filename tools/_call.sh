cd $GRAFT_REPO_ROOT
B="python bench.py --steps 2 --warmup 3 --streams 1 --views 2 --distinct 2 --cpu-views 0"
for K in k_occlusion_cull k_cull_compact k_light_prepass k_light_cells k_light_tiles; do
  ncu --set full --clock-control none --import-source on -k regex:$K -s 5 -c 1 -f -o gpurun_out/full_i_$K $B > gpurun_out/ncu_full_i_$K.log 2>&1
  tail -1 gpurun_out/ncu_full_i_$K.log
done
