cd $GRAFT_REPO_ROOT
bash tools/ab.sh c5 c7 2>&1 | grep -E "==|value|taa"
python bench.py --steps 5 --warmup 3 --cpu-views 0 > gpurun_out/ab_head.json 2> gpurun_out/ab_head.err; echo "== head"; python tools/show_bench.py gpurun_out/ab_head.json | grep -E "value|taa"
