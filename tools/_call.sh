cd $GRAFT_REPO_ROOT
(time python -m pytest tests -m gpu -x -q 2>&1 | tail -3) 2>&1 | grep -E "passed|failed|real"
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
(time python bench.py > gpurun_out/bench_j.json 2> gpurun_out/bench_j.err) 2>&1 | grep real; python tools/show_bench.py gpurun_out/bench_j.json | head -1
(time python bench.py --impl reference --steps 2 --warmup 0 > gpurun_out/bench_ref_j.json 2> gpurun_out/bench_ref_j.err) 2>&1 | grep real; head -c 150 gpurun_out/bench_ref_j.json
