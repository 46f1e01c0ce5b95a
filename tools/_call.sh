cd $GRAFT_REPO_ROOT
B="python bench.py --steps 2 --warmup 3 --streams 1 --views 2 --distinct 2 --cpu-views 0"
ncu --set full --clock-control none --import-source on -k regex:k_downsample_box_pipe -s 4 -c 1 -f -o gpurun_out/full_i_box_pipe $B > gpurun_out/ncu_full_i.log 2>&1
tail -1 gpurun_out/ncu_full_i.log
