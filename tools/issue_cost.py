#!/usr/bin/env python
"""Host-side cost of enqueueing one view (ashcuda_view_run through ctypes) vs the device time per view:
tells whether CUDA-graph replay of a view could matter.  Run on a GPU box: python tools/issue_cost.py"""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import bench  # noqa: E402
from ash_renderer_b200 import backend, types as T  # noqa: E402

objs, draws, lights = bench.make_scene()
d_objs = torch.from_numpy(objs.view(np.uint8)).cuda()
d_draws = torch.from_numpy(draws.view(np.uint8)).cuda()
d_lights = torch.from_numpy(lights.view(np.uint8)).cuda()
taa_pc = bench.taa_pc_default()
cam, depth, hdr = bench.make_view_inputs(0, 1024, objs)
d_depth, d_hdr = torch.from_numpy(depth).cuda(), torch.from_numpy(hdr.view(np.uint16)).cuda()
for streams, graph in ((4, 0), (4, 1), (1, 1)):
    ss = [torch.cuda.Stream() for _ in range(streams)]
    ctxs = []
    for v in range(16):
        c = backend.Context(bench.W, bench.H, max_objects=len(objs), max_lights=len(lights), hiz_levels=bench.HIZ_LEVELS, graph_replay=graph)
        c.set_stream(ss[v % streams].cuda_stream)
        c.set_objects(d_objs, d_draws)
        c.upload_lights(d_lights)
        ctxs.append(c)
    vp = bench.view_params(cam, len(objs), len(lights), taa_pc, 0)

    def step():
        for c in ctxs:
            c.set_depth(d_depth)
            c.set_color(d_hdr, None)
            c.view_run(vp)
    for _ in range(3):
        step()
    torch.cuda.synchronize()
    n = 20
    t0 = time.perf_counter()
    for _ in range(n):
        step()
    t_issue = time.perf_counter() - t0
    torch.cuda.synchronize()
    t_all = time.perf_counter() - t0
    l0 = ctxs[0].kernel_launches()
    step()
    print(f"streams={streams} graph_replay={graph}: host issue {t_issue / n / 16 * 1e6:.1f} us/view, end-to-end {t_all / n / 16 * 1e6:.1f} us/view, "
          f"{ctxs[0].kernel_launches() - l0} kernels/view")
    for c in ctxs:
        c.close()
