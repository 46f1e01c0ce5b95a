cd $GRAFT_REPO_ROOT
B="python bench.py --steps 2 --warmup 3 --streams 1 --views 2 --distinct 2 --cpu-views 0"
ncu --set full --clock-control none --import-source on -k regex:k_upsample_pair -s 5 -c 1 -f -o gpurun_out/full_g_k_upsample_pair_tm $B > gpurun_out/ncu_full_g_up.log 2>&1
tail -2 gpurun_out/ncu_full_g_up.log
