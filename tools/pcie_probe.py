"""Pinned host <-> device copy bandwidth of this box, on 1 GPU or on N GPUs AT ONCE (context for the end-to-end number).

    python tools/pcie_probe.py                                   # one GPU
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29511 tools/pcie_probe.py

Every rank copies 1 GiB buffers on its own GPU: host-to-device alone, device-to-host alone, then both directions at
once on two streams, all ranks between the same two barriers, timed on the device with CUDA events (max over ranks).
Rank 0 prints one JSON line: per-GPU and aggregate GB/s for the three cases.  The aggregate is the ceiling the host
fabric gives `bench.py`'s end-to-end pass, whatever the kernels do."""
import json
import os
import sys

import torch
import torch.distributed as dist

rank, world, local = (int(os.environ.get(k, d)) for k, d in (("RANK", 0), ("WORLD_SIZE", 1), ("LOCAL_RANK", 0)))
torch.cuda.set_device(local)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
try:                                                       # same CPU binding as bench.py
    import pynvml
    pynvml.nvmlInit()
    p = torch.cuda.get_device_properties(local)
    h_ = pynvml.nvmlDeviceGetHandleByPciBusId(f"{p.pci_domain_id:08x}:{p.pci_bus_id:02x}:{p.pci_device_id:02x}.0".encode())
    words = pynvml.nvmlDeviceGetCpuAffinity(h_, (os.cpu_count() + 63) // 64)
    cpus = {64 * i + b for i, w in enumerate(words) for b in range(64) if (w >> b) & 1} & os.sched_getaffinity(0)
    if cpus:
        os.sched_setaffinity(0, cpus)
except Exception:
    pass

n, reps = 1 << 30, 6
h1, h2 = torch.empty(n, dtype=torch.uint8).pin_memory(), torch.empty(n, dtype=torch.uint8).pin_memory()
d1, d2 = torch.empty(n, dtype=torch.uint8, device="cuda"), torch.empty(n, dtype=torch.uint8, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()


def up():
    with torch.cuda.stream(s1):
        d1.copy_(h1, non_blocking=True)


def down():
    with torch.cuda.stream(s2):
        h2.copy_(d2, non_blocking=True)


def timed(*fns):
    for f in fns:
        f()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    main = torch.cuda.current_stream()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(main)
    s1.wait_event(a); s2.wait_event(a)
    for _ in range(reps):
        for f in fns:
            f()
    for s in (s1, s2):
        e = torch.cuda.Event(); e.record(s); main.wait_event(e)
    b.record(main)
    torch.cuda.synchronize()
    ms = torch.tensor([a.elapsed_time(b)], device="cuda", dtype=torch.float64)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    return n * reps / (float(ms.item()) * 1e-3) / 1e9      # GB/s per GPU per direction, slowest rank


out = {"gpus": world, "bytes_per_copy": n, "h2d_alone_gbs_per_gpu": timed(up), "d2h_alone_gbs_per_gpu": timed(down),
       "duplex_gbs_per_gpu_each_way": timed(up, down)}
out["h2d_alone_gbs_all"] = out["h2d_alone_gbs_per_gpu"] * world
out["d2h_alone_gbs_all"] = out["d2h_alone_gbs_per_gpu"] * world
out["duplex_gbs_all_both_ways"] = out["duplex_gbs_per_gpu_each_way"] * world * 2
if rank == 0:
    print(json.dumps(out))
    print(json.dumps(out), file=sys.stderr)
if world > 1:
    dist.destroy_process_group()
