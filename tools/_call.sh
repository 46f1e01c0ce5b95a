cd $GRAFT_REPO_ROOT
python -m pytest tests -m gpu -x -q 2>&1 | tail -3
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/bench_h_2gpu.json 2> gpurun_out/bench_h_2gpu.err; tail -c 300 gpurun_out/bench_h_2gpu.err; python tools/show_bench.py gpurun_out/bench_h_2gpu.json | head -1
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --impl reference --gpus 2 --steps 1 --warmup 0 > gpurun_out/bench_ref_h_2gpu.json 2> gpurun_out/bench_ref_h_2gpu.err; head -c 300 gpurun_out/bench_ref_h_2gpu.json
