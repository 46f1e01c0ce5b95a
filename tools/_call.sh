cd $GRAFT_REPO_ROOT
cp ash_renderer_b200/libashcuda.so /tmp/keep.so
cp ash_renderer_b200/variants/libashcuda_th8.so ash_renderer_b200/libashcuda.so
python -m pytest tests -m gpu -x -q -k "taa or view_run" 2>&1 | tail -3
cp /tmp/keep.so ash_renderer_b200/libashcuda.so
python -m pytest tests -m gpu -x -q -k "taa or view_run" 2>&1 | tail -2
bash tools/ab.sh th8 2>&1 | grep -E "==|value|taa"
python bench.py --steps 5 --warmup 3 --cpu-views 0 > gpurun_out/ab_head.json 2> gpurun_out/ab_head.err; echo "== head"; python tools/show_bench.py gpurun_out/ab_head.json | grep -E "value|taa"
