cd $GRAFT_REPO_ROOT
python -m pytest tests -m gpu -x -q 2>&1 | tail -3
python bench.py --steps 5 --warmup 3 --cpu-views 0 > gpurun_out/ab_head.json 2> gpurun_out/ab_head.err || tail -3 gpurun_out/ab_head.err
echo "== head"; python tools/show_bench.py gpurun_out/ab_head.json | grep -E "value|tonemap"
bash tools/ab.sh prev 2>&1 | grep -E "==|value|tonemap"
