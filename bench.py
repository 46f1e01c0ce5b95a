#!/usr/bin/env python
"""bench.py — 4K views/sec of the per-frame compute path (Hi-Z + cull + light cull + TAA + bloom + tonemap).

One step = one frame of each of `--views` (32) independent camera views of the BASELINE.json config-5
scene (3840x2160, 100k instances, 8192 lights), every view with its own depth buffer, HDR colour
buffer and TAA history.  Views shard across GPUs (weak scaling: `--views` per GPU, no data-path
collective; one all_gather of the 32-byte per-view stats records at the end).  Beside the headline:
config 5 as written (a fixed batch of 1024 views sharded by views.shard_views: strong scaling), the
same step sustained for seconds, and a parity check of this very workload against the CPU oracle.

  value  : views/s with the per-view inputs already resident in HBM (CUDA events, max over ranks)
  e2e    : views/s through the host-buffer API: pinned host depth+HDR -> H2D -> view_run -> D2H of
           visibility, tile lists, 8-bit image and stats, every view, inside the timed region
  roofline : the dominant kernel class's algorithmic bytes per launch / its mean launch duration
             (CUDA events recorded by the library on the launching stream during the timed region)
  cpu_baseline : the CPU oracle (oracle/, OpenMP) on a bounded sample of the same workload
  --impl reference : the CPU oracle alone (the reference has no CPU implementation of this path
             and its Vulkan path cannot run here: see DESIGN.md), same metric/config.
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

W, H = 3840, 2160
N_OBJECTS_GRID = (50, 50, 40)      # 100 000 instances
N_LIGHTS = 8192
HIZ_LEVELS = 12                    # "full pyramid" (PIN-4)
METRIC, UNIT = "4k_views_per_sec", "views/s"
# dram__bytes_read.sum + dram__bytes_write.sum per launch, from the committed `ncu --set full` captures of this
# very workload (one launch each; re-captured whenever the kernel changes)
NCU_DRAM_SOURCE = "profiles/r02_p_full_k_taa_tol.md (132.72 MB read + 91.63 MB write, precision fast); exact mode: profiles/r01_j_full_k_taa_pair.md"
NCU_DRAM_BYTES = {("taa_threshold_fused", "fast"): 132.724736e6 + 91.633408e6, ("taa_threshold_fused", "exact"): 132.792064e6 + 93.815296e6}


def workload_name():
    return (f"cfg5: {W}x{H}, {np.prod(N_OBJECTS_GRID)} instances, {N_LIGHTS} lights, full Hi-Z ({HIZ_LEVELS} levels) + "
            "occlusion cull + 16x16 light cull + TAA + bloom(5) + ACES tonemap, 1 frame per view per step")


# ------------------------------------------------------------------ synthetic scene / views
def fibonacci_camera(i, n, rng):
    from ash_renderer_b200 import synth
    k = i + 0.5
    phi = np.arccos(1 - 2 * k / n)
    theta = np.pi * (1 + 5 ** 0.5) * k
    r = rng.uniform(30.0, 60.0)
    eye = r * np.array([np.cos(theta) * np.sin(phi), np.cos(phi) * 0.6, np.sin(theta) * np.sin(phi)])
    target = rng.uniform(-5.0, 5.0, 3)
    return synth.Camera(eye, target, W, H, near=0.5, far=200.0)


def make_scene():
    from ash_renderer_b200 import synth
    objs, draws = synth.make_objects(N_OBJECTS_GRID, spacing=2.0, seed=5, jitter=0.4)
    lights = synth.make_lights(N_LIGHTS, extent=60.0, rmin=1.0, rmax=6.0, seed=5)
    return objs, draws, lights


def make_view_inputs(view_id, n_total, objs):
    """(camera, depth fp32 HxW, hdr fp16 HxWx4) for one view; deterministic in view_id."""
    from ash_renderer_b200 import synth
    rng = np.random.default_rng(1000 + view_id)
    cam = fibonacci_camera(view_id, n_total, rng)
    depth = synth.splat_depth(objs[:: 8], cam, shrink=0.7)            # coherent occluders from 1/8 of the instances
    hdr = synth.HdrSequence(W, H, seed=5 + view_id, n_discs=120).frame(0.0, 0.0)
    return cam, depth, hdr


def view_params(cam, n_objs, n_lights, taa_pc, view_id):
    from ash_renderer_b200 import types as T
    vp = np.zeros(1, T.VIEW_PARAMS)
    vp["cull"] = cam.cull_pc(n_objs, HIZ_LEVELS)
    vp["light"] = cam.light_pc(n_lights)
    vp["camera"] = cam.camera_data()
    vp["taa"] = taa_pc
    vp["bloom"]["threshold"], vp["bloom"]["upsample_intensity"], vp["bloom"]["mip_count"] = 0.8, 1.0, 5
    vp["post"]["exposure"], vp["post"]["gamma"], vp["post"]["bloom_intensity"] = 1.0, 2.2, 0.5
    # (ASH_BENCH_STAGES: experiments only — e.g. 0x38 times the post chain alone; the reported workload is always STAGE_ALL)
    vp["view_id"], vp["stages"] = view_id, int(os.environ.get("ASH_BENCH_STAGES", "0"), 0) or T.STAGE_ALL
    return vp


# ------------------------------------------------------------------ algorithmic bytes
def algorithmic_bytes(kclass, n_objs, n_lights, st):
    from ash_renderer_b200 import types as T  # noqa: F401
    P = W * H
    c = 8
    mips = []
    w, h = W, H
    while max(w, h) > 1:
        w, h = max(1, w // 2), max(1, h // 2)
        mips.append(w * h)
    tiles = ((W + 15) // 16) * ((H + 15) // 16)
    bl = [P]
    w, h = W, H
    for _ in range(5):
        w, h = max(1, w // 2), max(1, h // 2)
        bl.append(w * h)
    V, cand, sum_l = int(st["visible_objects"]), int(st["candidate_lights"]), int(st["sum_tile_lights"])
    TAIL = 1 << 20
    mean = lambda v: float(sum(v)) / max(len(v), 1)
    table = {
        0: 4 * P + 4 * sum(mips) + 8 * tiles,                                # depth read once, mips + tile min/max written
        1: 112 * n_objs + 4 * n_objs + 4 * n_objs + n_objs // 8,             # objects, visibility, one Hi-Z texel, bit mask
        2: n_objs // 8 + 4 * V + 40 * V + 4,                                 # mask, draw index, command in+out
        3: 64 * n_lights + 208 + 20 * cand,
        4: 8 * tiles + 20 * cand + 4 * (tiles + sum_l),
        5: 3 * c * P,
        6: 2 * c * P,
        # bloom levels by size class (ASH_KERNEL_BLOOM_TAIL: destination below 2^20 texels): mean per launch of each class
        7: mean([c * (bl[k - 1] + bl[k]) for k in range(1, 6) if bl[k] >= TAIL]),
        8: mean([c * (bl[k + 1] + 2 * bl[k]) for k in range(1, 5) if bl[k] >= TAIL]),
        15: mean([c * (bl[k - 1] + bl[k]) for k in range(1, 6) if bl[k] < TAIL] + [c * (bl[k + 1] + 2 * bl[k]) for k in range(1, 5) if bl[k] < TAIL]),
        9: 2 * c * P + 4 * P,
        10: 4 * c * P,                                                       # fused TAA + threshold: cur, hist in; out, T out
        11: c * (bl[1] + 2 * P) + 4 * P,                                     # fused last upsample + tonemap
        # row f-2 (--instancing): visibility + batch id + mesh id, position of visible objects, selection out,
        # kept instances in + out, one draw per bucket
        12: 12 * n_objs + 16 * V + 16 * n_objs + 192 * int(st.get("kept_instances", 0)) + 20 * int(st.get("buckets", 0)),
        # row f-1 (--forward): objects in, sphere + rectangle out, key clear, key resolve (8 B in, 4 B depth out);
        # the splat's atomics hit L2.  Shade: key in, RGBA16F out, the tile lists.
        13: 144 * n_objs + 8 * P + 12 * P,
        14: 8 * P + c * P + 4 * (tiles + sum_l),
    }
    return float(table[kclass])


def compulsory_bytes(n_objs, n_lights, st):
    """SURVEY.md 8.4, the headline definition: every INPUT of the view once and every output the API exposes once
    (depth, all Hi-Z mips, objects, lights, visibility, compacted draws, tile lists, current + history colour, the
    resolve = next history, LDR8) — whatever the fused pipeline stages in between costs nothing here."""
    P, c = W * H, 8
    mips, w, h = [], W, H
    while max(w, h) > 1:
        w, h = max(1, w // 2), max(1, h // 2)
        mips.append(w * h)
    tiles = ((W + 15) // 16) * ((H + 15) // 16)
    V, sum_l = int(st["visible_objects"]), int(st["sum_tile_lights"])
    return float(4 * P + 4 * sum(mips) + 112 * n_objs + 64 * n_lights + 208 + 4 * n_objs + 20 * V + 4 + 4 * (tiles + sum_l)
                 + 3 * c * P + 4 * P)


KERNEL_NAMES = ["depth_pyramid", "occlusion_cull", "cull_compact", "light_prepass", "light_tiles", "taa_resolve",
                "bloom_threshold", "bloom_downsample", "bloom_upsample", "tonemap", "taa_threshold_fused",
                "upsample_tonemap_fused", "instancing", "forward_depth", "forward_shade", "bloom_tail_small_levels"]


# ------------------------------------------------------------------ clocks sampling (B200_PROFILING.md)
class ClockSampler:
    Q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, mx, reasons, power = [], [], set(), []
        for r in self.rows:
            if len(r) < 7:
                continue
            try:
                sm.append(float(r[0])); mx.append(float(r[1])); power.append(float(r[2]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w_max": max(power) if power else None, "samples": len(sm), "reasons": sorted(reasons)}


# ------------------------------------------------------------------ CPU oracle arm
def oracle_view(oracle, objs, draws, lights, cam, depth, hdr, hist, taa_pc):
    from ash_renderer_b200 import types as T
    mips, flat = oracle.hiz_build(depth)
    vis, dout, cnt = oracle.occlusion_cull(objs, draws, cam.cull_pc(len(objs), HIZ_LEVELS), depth, flat)
    td, _, _ = oracle.light_cull(lights, cam.light_pc(len(lights)), cam.camera_data(), depth)
    taa = oracle.taa_resolve(hdr, hist, None, taa_pc)
    U, _ = oracle.bloom_chain(taa, 0.8, 5)
    pc = np.zeros(1, T.POST_PC)
    pc["exposure"], pc["gamma"], pc["bloom_intensity"] = 1.0, 2.2, 0.5
    ldr = oracle.tonemap(taa, U[0], pc)
    return vis, cnt, td, taa, ldr


def cpu_oracle_rate(objs, draws, lights, views, taa_pc, max_views, budget_s):
    """views/s of the OpenMP oracle on up to `max_views` full 4K views (bounded by `budget_s`), and what it computed
    (kept: the CUDA results of the same views are checked against it, `parity_check`)."""
    from oracle import oracle
    oracle.build()
    if max_views <= 0:
        return None, oracle.num_threads(), 0, []
    times, results = [], []
    t_all = time.perf_counter()
    for i in range(max_views):
        cam, depth, hdr = views[i % len(views)]
        t0 = time.perf_counter()
        results.append(oracle_view(oracle, objs, draws, lights, cam, depth, hdr, hdr, taa_pc))
        times.append(time.perf_counter() - t0)
        if time.perf_counter() - t_all > budget_s:
            break
    return len(times) / sum(times), oracle.num_threads(), len(times), results


def parity_check(slots, oracle_results, fast):
    """The benchmark workload itself against the oracle: slot i re-runs view i from a fresh TAA history (first frame:
    history = current colour, as the oracle arm ran it) and every output is compared — visibility, visible count and tile
    light lists bit-exact; the TAA resolve bit-exact (exact mode) or within 1 fp16 ulp (fast mode); LDR8 within 1 LSB."""
    n = 0
    for s, (vis, cnt, td, taa, ldr) in zip(slots, oracle_results):
        c = s["ctx"]
        c.taa_reset_history()
        c.set_depth(s["d_depth"])
        c.set_color(s["d_hdr"], None)
        c.view_run(s["vp"])
        got_td = c.tile_data()
        assert np.array_equal(c.visibility(), vis) and c.visible_count() == cnt, "bench workload: visibility differs from the oracle"
        assert np.array_equal(got_td[:, 0], td[:, 0]), "bench workload: tile light counts differ from the oracle"
        m = np.arange(td.shape[1])[None, :] < (1 + td[:, :1])
        assert np.array_equal(got_td[m], td[m]), "bench workload: tile light lists differ from the oracle"
        g = c.taa_out().view(np.uint16).astype(np.int32)
        r = taa.view(np.uint16).astype(np.int32)
        d = np.abs(np.where(g & 0x8000, -(g & 0x7FFF), g) - np.where(r & 0x8000, -(r & 0x7FFF), r))
        assert d.max() <= (1 if fast else 0), f"bench workload: TAA differs from the oracle by {int(d.max())} fp16 ulp"
        assert np.abs(c.ldr8().astype(np.int16) - ldr.astype(np.int16)).max() <= 1, "bench workload: LDR8 differs by more than 1 LSB"
        n += 1
    return n


def taa_pc_default():
    from ash_renderer_b200.host import TemporalAA
    t = TemporalAA()
    t.begin_frame(); t.begin_frame()
    return t.push_constants(W, H)


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    # torchrun exports OMP_NUM_THREADS=1; the CPU arm is meant to use every host core it can
    os.environ["OMP_NUM_THREADS"] = str(os.cpu_count() or 1)
    objs, draws, lights = make_scene()
    n_views = 2
    views = [make_view_inputs(v, 1024, objs) for v in range(n_views)]
    pc = taa_pc_default()
    from oracle import oracle
    oracle.build()
    for i in range(args.warmup):
        if i >= 1:
            break                                          # one warm-up view is enough to page everything in
        cam, depth, hdr = views[0]
        oracle_view(oracle, objs, draws, lights, cam, depth, hdr, hdr, pc)
    t0 = time.perf_counter()
    for k in range(args.steps):
        cam, depth, hdr = views[k % n_views]
        oracle_view(oracle, objs, draws, lights, cam, depth, hdr, hdr, pc)
    dt = time.perf_counter() - t0
    v = args.steps / dt
    sample = f"{args.steps} steps x 1 full 4K view each (bounded sample of the {args.views}-view step)"
    emit_json({
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": min(args.warmup, 1), "ms_per_step": 1e3 * dt / args.steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": workload_name(), "views_per_step": 1, "note": "CPU oracle port of the reference GLSL; "
                   "the reference has no CPU implementation and its Vulkan path is unavailable on this box"},
        "cpu_baseline": {"value": v, "unit": UNIT, "cores": oracle.num_threads(), "kind": "port", "sample": sample},
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    })


def bind_cpu_near_gpu(torch, local):
    """Pin this rank's threads (and, by first touch, its pinned staging buffers) to the CPUs NVML reports as
    local to its GPU: the end-to-end pass moves 166 MB per view through host memory."""
    try:
        import pynvml
        pynvml.nvmlInit()
        p = torch.cuda.get_device_properties(local)
        h = pynvml.nvmlDeviceGetHandleByPciBusId(f"{p.pci_domain_id:08x}:{p.pci_bus_id:02x}:{p.pci_device_id:02x}.0".encode())
        words = pynvml.nvmlDeviceGetCpuAffinity(h, (os.cpu_count() + 63) // 64)
        cpus = {64 * i + b for i, w in enumerate(words) for b in range(64) if (w >> b) & 1} & os.sched_getaffinity(0)
        if cpus:
            os.sched_setaffinity(0, cpus)
        return len(cpus)
    except Exception as e:          # affinity is an optimisation only
        print(f"[bench] cpu affinity not set: {e}", file=sys.stderr)
        return 0


# ------------------------------------------------------------------ CUDA arm
def run_cuda(args):
    import torch
    import torch.distributed as dist
    from ash_renderer_b200 import backend, synth, types as T

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    ncpu_local = bind_cpu_near_gpu(torch, local)
    print(f"[bench] rank {rank}: {ncpu_local} CPUs local to GPU {local}", file=sys.stderr)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    B = args.views
    objs, draws, lights = make_scene()
    n_objs, n_lights = len(objs), len(lights)
    total_views = B * world
    distinct = min(B, args.distinct)
    taa_pc = taa_pc_default()

    # scene: replicated once per GPU, shared (borrowed) by every view slot
    d_objs = torch.from_numpy(objs.view(np.uint8)).cuda()
    d_draws = torch.from_numpy(draws.view(np.uint8)).cuda()
    d_lights = torch.from_numpy(lights.view(np.uint8)).cuda()

    inst_scene = None
    if args.instancing:
        from ash_renderer_b200 import host as H_
        meshes = synth.make_lod_meshes(6, seed=7)
        inst, batch, mesh = synth.make_instances(objs, 16, 6, seed=8)
        inst_scene = {"meshes": meshes, "inst": inst, "batch": batch, "mesh": mesh, "cfg": H_.LodConfig().record(use_visibility=True),
                      "d_inst": torch.from_numpy(inst.view(np.uint8)).cuda(), "d_batch": torch.from_numpy(batch.view(np.int32)).cuda(),
                      "d_mesh": torch.from_numpy(mesh.view(np.int32)).cuda(), "d_meshes": torch.from_numpy(meshes.view(np.uint8)).cuda()}

    fwd_params = synth.forward_params(near=0.5, flags=T.FWD_TILE_LIGHTS) if args.forward else None
    # `distinct` inputs are generated on the host (seconds each); every slot then gets its OWN variant of one of them,
    # shifted horizontally by a slot-specific amount on the device, so that no two views of a GPU share an input
    # (SURVEY 8.4 cfg-5 allows 32 distinct buffers cycled over the 1024 views; B >= 32 gives that)
    host_views = [make_view_inputs(rank * B + v, max(total_views, 1024), objs) for v in range(distinct)]
    streams = [torch.cuda.Stream() for _ in range(args.streams)]
    slots = []
    for v in range(B):
        cam, depth, hdr = host_views[v % distinct]
        c = backend.Context(W, H, device=local, max_objects=n_objs, max_lights=n_lights, hiz_levels=HIZ_LEVELS,
                            graph_replay=args.graph_replay, precision=T.PRECISION_FAST if args.precision == "fast" else T.PRECISION_EXACT)
        c.set_stream(streams[v % args.streams].cuda_stream)
        c.set_objects(d_objs, d_draws)
        c.upload_lights(d_lights)
        vparams = view_params(cam, n_objs, n_lights, taa_pc, rank * B + v)
        if inst_scene:
            c.set_instances(inst_scene["d_inst"], inst_scene["d_batch"], inst_scene["d_mesh"])
            c.set_lod_meshes(inst_scene["d_meshes"])
            c.set_lod_config(inst_scene["cfg"])
            vparams["stages"] |= T.STAGE_INSTANCING
        if args.forward:
            c.set_forward_params(fwd_params)
            vparams["stages"] |= T.STAGE_FORWARD_DEPTH | T.STAGE_FORWARD_SHADE
            if args.two_phase:
                vparams["stages"] |= T.STAGE_TWO_PHASE
        shift = (v // distinct) * 211                       # 0 for the first `distinct` slots (those the oracle arm also runs)
        d_depth = torch.roll(torch.from_numpy(depth).cuda(), shift, dims=1).contiguous()
        d_hdr = torch.roll(torch.from_numpy(hdr.view(np.int16)).cuda(), shift, dims=1).contiguous()     # (roll has no uint16 kernel)
        slot = {"ctx": c, "vp": vparams, "d_depth": d_depth, "d_hdr": d_hdr,
                "h_depth": d_depth.cpu().pin_memory(), "h_hdr": d_hdr.cpu().pin_memory()}
        tiles = ((W + 15) // 16) * ((H + 15) // 16)
        slot["o_vis"] = torch.empty(n_objs, dtype=torch.int32).pin_memory()
        slot["o_off"] = torch.empty(tiles + 1, dtype=torch.int32).pin_memory()      # tile light lists in CSR form:
        slot["o_tl"] = torch.empty(tiles * 16, dtype=torch.int32).pin_memory()      # offsets + up to 16 lights per tile on average
        slot["o_ldr"] = torch.empty(W * H * 4, dtype=torch.uint8).pin_memory()
        slot["o_st"] = torch.empty(8, dtype=torch.int32).pin_memory()
        slot["o_cnt"] = torch.empty(1, dtype=torch.int32).pin_memory()
        slots.append(slot)
    torch.cuda.synchronize()

    main = torch.cuda.current_stream()

    def fence_start():
        ev = torch.cuda.Event(enable_timing=True)
        ev.record(main)
        for s in streams:
            s.wait_event(ev)
        return ev

    def fence_stop():
        for s in streams:
            e = torch.cuda.Event()
            e.record(s)
            main.wait_event(e)
        ev = torch.cuda.Event(enable_timing=True)
        ev.record(main)
        return ev

    d_vel = None
    if args.velocity:
        d_vel = torch.from_numpy(synth.random_velocity(W, H, seed=11, still_frac=0.3, max_px=4.0).view(np.uint16)).cuda()

    def step_resident():
        for s in slots:
            c = s["ctx"]
            if not args.forward:                       # --forward: depth and HDR colour are made inside view_run
                c.set_depth(s["d_depth"])
                c.set_color(s["d_hdr"], d_vel)
            c.view_run(s["vp"])

    def step_e2e():
        for s in slots:
            c = s["ctx"]
            if not args.forward:                       # --forward: only the parameter block crosses the host boundary
                c.upload_view(s["h_depth"], s["h_hdr"])
            c.view_run(s["vp"])
            c.download_view_packed(s["o_vis"], s["o_off"], s["o_tl"], s["o_ldr"], s["o_st"], s["o_cnt"])

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(step_fn, steps, warmup):
        for _ in range(warmup):
            step_fn()
        barrier()
        a = fence_start()
        for _ in range(steps):
            step_fn()
        b = fence_stop()
        barrier()
        ms = a.elapsed_time(b)
        if world > 1:
            t = torch.tensor([ms], device="cuda", dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms

    # ---- device-resident throughput (the `value`): every stream busy, front end forked per view
    for s in slots:
        s["ctx"].profile_kernels(0)
    timed(step_resident, 1, max(args.warmup, 3) - 1)       # warm-up (also fills every TAA history)
    launches0 = sum(s["ctx"].kernel_launches() for s in slots)
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
        time.sleep(0.3)
    ms = timed(step_resident, args.steps, 1)
    clocks = sampler.stop() if rank == 0 else None
    launches_per_step = (sum(s["ctx"].kernel_launches() for s in slots) - launches0) / (args.steps + 1)
    value = total_views * args.steps / (ms * 1e-3)

    # ---- sustained: the same step looped for >= 3 s (the timed region above is tens of milliseconds), own clock sample
    sustained = None
    if args.sustain_s > 0:
        n_sus = max(args.steps, int(np.ceil(args.sustain_s * 1e3 / (ms / args.steps))))
        sampler2 = ClockSampler(local)
        if rank == 0:
            sampler2.start()
        ms_sus = timed(step_resident, n_sus, 0)
        clocks_sus = sampler2.stop() if rank == 0 else None
        sustained = {"views_per_s": total_views * n_sus / (ms_sus * 1e-3), "seconds": ms_sus * 1e-3, "steps": n_sus, "clocks": clocks_sus}

    # ---- BASELINE config 5 as written: a FIXED batch of 1024 independent camera views, sharded over the GPUs by
    #      views.shard_views (strong scaling).  Every view has its own camera; slot v's depth / colour input serves the
    #      views v, v + B, ... of this GPU's shard (B distinct inputs per GPU).
    from ash_renderer_b200 import views as V
    cfg5 = None
    if args.total_views > 0 and not args.forward:
        mine = V.shard_views(args.total_views, world, rank)
        rng5 = np.random.default_rng(7)
        cams = [fibonacci_camera(v, args.total_views, rng5) for v in range(args.total_views)]   # same sequence on every rank
        vps = []
        for i, v in enumerate(mine):
            vp5 = slots[i % B]["vp"].copy()
            vp5["cull"], vp5["light"], vp5["camera"], vp5["view_id"] = cams[v].cull_pc(n_objs, HIZ_LEVELS), cams[v].light_pc(n_lights), cams[v].camera_data(), v
            vps.append(vp5)

        def run_shard():
            for i in range(len(mine)):
                sl = slots[i % B]
                sl["ctx"].set_depth(sl["d_depth"])
                sl["ctx"].set_color(sl["d_hdr"], d_vel)
                sl["ctx"].view_run(vps[i])
        ms5 = timed(run_shard, 1, 1)
        cfg5 = {"total_views": args.total_views, "views_this_rank": len(mine), "sharding": "views.shard_views (block)",
                "distinct_inputs_per_gpu": B, "scaling": "strong", "ms": ms5, "views_per_s": args.total_views / (ms5 * 1e-3)}

    # ---- per-kernel device time: the same steps again in the same process, serialised (one stream,
    #      no fork) so that each CUDA-event pair recorded by the library brackets exactly one kernel
    for s in slots:
        s["ctx"].set_stream(streams[0].cuda_stream)
        s["ctx"].profile_kernels((1 << T.KERNEL_COUNT) - 1)
        for k in range(T.KERNEL_COUNT):
            s["ctx"].kernel_time_ms(k)                     # reset accumulators
    prof_ms = timed(step_resident, args.steps, 1)
    ktime = np.zeros(T.KERNEL_COUNT)
    kcount = np.zeros(T.KERNEL_COUNT)
    for s in slots:
        for k in range(T.KERNEL_COUNT):
            t, n = s["ctx"].kernel_time_ms(k)
            ktime[k] += t
            kcount[k] += n
        s["ctx"].profile_kernels(0)
    for v, s in enumerate(slots):
        s["ctx"].set_stream(streams[v % args.streams].cuda_stream)
    stats = [s["ctx"].stats() for s in slots]
    st_mean = {f: float(np.mean([st[f] for st in stats])) for f in T.VIEW_STATS.names}
    widened = None
    if inst_scene:
        ist = [s["ctx"].instancing_stats() for s in slots]
        st_mean["kept_instances"] = float(np.mean([x["total_instances"] for x in ist]))
        st_mean["buckets"] = 16 * 8
        widened = {"row": "f-2: LodManager::select_lod + InstancingManager::add_instance/finalize on the device",
                   "instances": n_objs, "batches": 16, "lod_meshes": 6,
                   "kept_instances_mean": st_mean["kept_instances"],
                   "objects_culled_by_lod_mean": float(np.mean([x["objects_culled"] for x in ist])),
                   "draw_calls_mean": float(np.mean([x["batch_count"] for x in ist])),
                   "triangle_reduction_mean": float(np.mean([1.0 - x["triangles_after"] / max(int(x["triangles_before"]), 1) for x in ist]))}
        if world == 1 and args.cpu_views > 0:     # the serial CPU managers on the same scene (oracle restatement)
            from oracle import oracle as O
            cam0 = host_views[0][0]
            vis0 = slots[0]["ctx"].visibility()
            t0 = time.perf_counter()
            O.instancing(inst_scene["inst"], inst_scene["batch"], inst_scene["mesh"], inst_scene["meshes"],
                         synth.instancing_params(cam0, n_objs), inst_scene["cfg"], 16, vis0)
            widened["cpu_ms_per_view"] = (time.perf_counter() - t0) * 1e3

    if args.forward:
        keys0 = slots[0]["ctx"].forward_keys()
        widened = dict(widened or {})
        widened.update({"row_f1": "depth + HDR colour produced on the device: occluder-sphere ray cast + tiled-forward Cook-Torrance shading "
                                  "(frag.frag BRDF) over the tile light lists; no per-view host upload",
                        "two_phase_culling": bool(args.two_phase),
                        "culling_stats": {k: int(slots[0]["ctx"].culling_stats()[k]) for k in T.CULLING_STATS.names},
                        "occluders": n_objs, "covered_fraction": float(((keys0 & np.uint64(0xFFFFFFFF)) != np.uint64(0xFFFFFFFF)).mean())})
        if world == 1 and args.cpu_views > 0:     # the CPU oracle of the same two stages on view 0 (serial splat, OpenMP shade)
            from oracle import oracle as O
            cam0 = host_views[0][0]
            t0 = time.perf_counter()
            k0, d0 = O.forward_depth(objs, cam0.camera_data(), fwd_params, W, H)
            t1 = time.perf_counter()
            O.forward_shade(k0, objs, lights, slots[0]["ctx"].tile_data(), cam0.camera_data(), fwd_params)
            widened["cpu_forward_depth_ms"] = (t1 - t0) * 1e3
            widened["cpu_forward_shade_ms"] = (time.perf_counter() - t1) * 1e3
            if not args.two_phase:            # (two-phase draws only the visible objects: a different, history-dependent image)
                widened["forward_depth_keys_equal_oracle"] = bool(np.array_equal(k0, keys0))

    # ---- latency of ONE view with nothing else in flight (what a single-camera frame loop sees): 20 views, one at a
    #      time on one slot, device time per view (informational: `value` is throughput with views overlapped)
    def one_view():
        c = slots[0]["ctx"]
        if not args.forward:
            c.set_depth(slots[0]["d_depth"])
            c.set_color(slots[0]["d_hdr"], d_vel)
        c.view_run(slots[0]["vp"])
    for _ in range(3):
        one_view()
    torch.cuda.synchronize()
    lat = []
    for _ in range(20):
        a, b = fence_start(), None
        one_view()
        b = fence_stop()
        torch.cuda.synchronize()
        lat.append(a.elapsed_time(b))
    single_view_ms = float(np.median(lat))

    # ---- end-to-end through host buffers
    e2e_ms = timed(step_e2e, args.steps, 2)
    e2e_value = total_views * args.steps / (e2e_ms * 1e-3)
    # each rank's own copy rates in that pass (its own device time; the headline uses the max over ranks)
    e2e_local = None
    if world > 1:
        tl = torch.zeros(world, device="cuda", dtype=torch.float64)
        a_, b_ = fence_start(), None
        step_e2e()
        b_ = fence_stop()
        torch.cuda.synchronize()
        tl[rank] = a_.elapsed_time(b_)
        dist.all_reduce(tl)
        e2e_local = [float(x) for x in tl.tolist()]
    tiles = ((W + 15) // 16) * ((H + 15) // 16)
    h2d = B * (W * H * 4 + W * H * 8 + T.VIEW_PARAMS.itemsize)   # depth + HDR + the per-view parameter block
    if args.forward:
        h2d = B * T.VIEW_PARAMS.itemsize
    d2h = B * (n_objs * 4 + (tiles + 1) * 4 + tiles * 16 * 4 + W * H * 4 + 32 + 4)   # visibility, packed tile lists, LDR8, stats, count
    # sanity: the e2e outputs equal the resident path's
    chk = slots[0]
    assert int(chk["o_cnt"][0]) == int(stats[0]["visible_objects"]), "e2e result differs from resident result"
    off = chk["o_off"].numpy().view(np.uint32)
    table = chk["ctx"].tile_data()
    assert int(off[-1]) == int(table[:, 0].sum()) <= tiles * 16, "packed tile lists truncated or inconsistent"
    t_probe = int(np.argmax(table[:, 0]))
    assert np.array_equal(chk["o_tl"].numpy().view(np.uint32)[off[t_probe]:off[t_probe + 1]], table[t_probe, 1:1 + table[t_probe, 0]])

    # ---- stats all_gather (the only inter-GPU exchange; 32 B per view)
    st_all = V.stats_to_words(V.all_gather_stats(np.array(stats, T.VIEW_STATS), total_views, world, rank,
                                                 device="cuda" if world > 1 else None))
    assert len(st_all) == total_views, "stats all_gather lost views"
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except OSError:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    peak_src = "MEASURED_PEAKS.json hbm_gbs (measured copy)" if "hbm_gbs" in peaks else "fallback 6650 GB/s (B200_PROFILING.md)"
    kernels = {}
    for k in range(T.KERNEL_COUNT):
        if kcount[k] == 0:
            continue
        avg_ms = ktime[k] / kcount[k]
        ab = algorithmic_bytes(k, n_objs, n_lights, st_mean)
        kernels[KERNEL_NAMES[k]] = {"launches_per_view": kcount[k] / (B * (args.steps + 1)), "avg_ms": avg_ms,
                                    "alg_bytes_per_launch": ab, "gbs": ab / (avg_ms * 1e-3) / 1e9,
                                    "share_of_kernel_time": ktime[k] / ktime.sum()}
    dom = max(range(T.KERNEL_COUNT), key=lambda k: ktime[k])
    dom_k = kernels[KERNEL_NAMES[dom]]
    # DRAM bytes per launch of that kernel (dram__bytes_read.sum + dram__bytes_write.sum) from the committed
    # `ncu --set full` capture of this workload; null for any other workload or kernel.
    traffic = NCU_DRAM_BYTES.get((KERNEL_NAMES[dom], args.precision))          # the workload (cfg5, 4K) is fixed in this file
    # whole-view roofline: compulsory bytes of the staged path / measured time per view
    view_bytes = sum(algorithmic_bytes(k, n_objs, n_lights, st_mean) * kcount[k] / (B * (args.steps + 1))
                     for k in range(T.KERNEL_COUNT) if kcount[k])
    per_view_s = ms * 1e-3 / (B * args.steps)

    # ---- CPU baseline (bounded sample: up to 2 full 4K views, ~10-30 s)
    # (rank 0 at N=1 only: under torchrun the other ranks would wait at exit and OMP is pinned to 1 thread)
    cpu_v, cpu_cores, cpu_n, oracle_results = cpu_oracle_rate(objs, draws, lights, host_views, taa_pc,
                                                              max_views=args.cpu_views if world == 1 else 0, budget_s=20.0)
    plain = not (args.forward or args.velocity or inst_scene)
    parity_checked = parity_check(slots, oracle_results, args.precision == "fast") if plain else 0
    comp_bytes = compulsory_bytes(n_objs, n_lights, st_mean)

    out = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
        "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": {"workload": workload_name() + (" + LOD selection / instance batching (--instancing)" if inst_scene else "")
                   + (" with depth + HDR colour produced on the device by the compute forward pass (--forward, row f-1)" if args.forward else "")
                   + (" + two-phase occlusion culling (--two-phase, row f-3)" if args.two_phase else "")
                   + (" + per-pixel velocity buffer in the TAA (--velocity; resident passes only)" if args.velocity else ""),
                   "views_per_gpu_per_step": B, "distinct_view_inputs_per_gpu": distinct,
                   "streams": args.streams, "concurrent_frontend": True,
                   "graph_replay": bool(args.graph_replay), "precision": args.precision, "single_view_latency_ms": single_view_ms,
                   "kernel_timing": f"same {args.steps} steps repeated serialised (1 stream, no fork) with CUDA events around each launch: "
                                    f"{prof_ms / args.steps:.3f} ms/step", "l2": f"inputs larger than L2: {B} views x 99.5 MB of depth+HDR per step per GPU",
                   "sustained_views_per_s": sustained["views_per_s"] if sustained else None, "sustained": sustained,
                   "cfg5_1024_views": cfg5,
                   "parity_checked_views": parity_checked,
                   "parity_check": "the CUDA results of the views the CPU arm ran (this very workload) compared with the oracle's: "
                                   "visibility, visible count, tile light lists bit-exact; TAA "
                                   + ("within 1 fp16 ulp" if args.precision == "fast" else "bit-exact") + "; LDR8 within 1 LSB",
                   "visible_fraction_mean": st_mean["visible_objects"] / n_objs,
                   "candidate_lights_mean": st_mean["candidate_lights"],
                   "lights_per_tile_mean": st_mean["sum_tile_lights"] / tiles, "lights_per_tile_max": int(st_all[:, 5].max())},
        "clocks": clocks,
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                "ms_per_step": e2e_ms / args.steps,
                "h2d_gbs_per_gpu": h2d / (e2e_ms / args.steps * 1e-3) / 1e9, "d2h_gbs_per_gpu": d2h / (e2e_ms / args.steps * 1e-3) / 1e9,
                "h2d_gbs_all_gpus": world * h2d / (e2e_ms / args.steps * 1e-3) / 1e9,
                "one_step_ms_by_rank": e2e_local},
        "gpu_launches": int(round(launches_per_step * args.steps)),
        "roofline": {"bound": "hbm", "kernel": KERNEL_NAMES[dom], "achieved": dom_k["gbs"], "peak": peak, "unit": "GB/s",
                     "frac": dom_k["gbs"] / peak, "traffic": traffic, "traffic_source": NCU_DRAM_SOURCE if traffic else None,
                     "peak_source": peak_src, "frac_of_nominal_8000_gbs": dom_k["gbs"] / 8000.0,
                     "whole_view": {"staged": {"definition": "sum over the launches of one view of each kernel's algorithmic bytes "
                                                             "(intermediates between kernels counted once written, once read)",
                                               "bytes_per_view": view_bytes, "gbs": view_bytes / per_view_s / 1e9,
                                               "frac": view_bytes / per_view_s / 1e9 / peak},
                                    "compulsory": {"definition": "SURVEY 8.4 headline: the view's inputs once + the outputs the API exposes "
                                                                 "once (depth, mips, objects, lights, visibility, draws, tile lists, current, "
                                                                 "history, resolve, LDR8)",
                                                   "bytes_per_view": comp_bytes, "gbs": comp_bytes / per_view_s / 1e9,
                                                   "frac": comp_bytes / per_view_s / 1e9 / peak},
                                    "alg_bytes_per_view_staged": view_bytes, "gbs": view_bytes / per_view_s / 1e9,
                                    "frac": view_bytes / per_view_s / 1e9 / peak,
                                    "frac_of_nominal_8000_gbs": view_bytes / per_view_s / 1e9 / 8000.0}},
        "kernels": kernels,
        "widened": widened,
        "cpu_baseline": {"value": cpu_v, "unit": UNIT, "cores": cpu_cores, "kind": "port",
                         "sample": f"{cpu_n} full 4K view(s) of the same workload through the OpenMP CPU oracle"},
    }
    emit_json(out)
    if world > 1:
        dist.destroy_process_group()


_JSON_FD = None


def emit_json(obj):
    """The one JSON line of the contract goes to the process's ORIGINAL stdout; everything else any library
    prints (NCCL's version banner, torchrun chatter) was redirected to stderr at start-up."""
    line = (json.dumps(obj) + "\n").encode()
    if _JSON_FD is None:
        sys.stdout.write(line.decode())
        sys.stdout.flush()
    else:
        os.write(_JSON_FD, line)


def main():
    global _JSON_FD
    sys.stdout.flush()
    _JSON_FD = os.dup(1)
    os.dup2(2, 1)                      # fd 1 -> stderr for the rest of the run
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=8)       # 8 steps x 32 views = 256 views per GPU
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="cuda", choices=["cuda", "reference"])
    ap.add_argument("--views", type=int, default=32, help="views per GPU per step (each with its own depth / colour input and TAA history)")
    ap.add_argument("--sustain-s", type=float, default=3.0, help="also loop the resident step for at least this many seconds (0: skip)")
    ap.add_argument("--total-views", type=int, default=1024,
                    help="BASELINE config 5 as written: a fixed batch of this many camera views sharded over the GPUs (strong scaling), "
                         "reported under config.cfg5_1024_views beside the weak-scaling headline (0: skip)")
    ap.add_argument("--distinct", type=int, default=4, help="distinct synthetic depth/HDR inputs generated per GPU (host setup cost)")
    ap.add_argument("--streams", type=int, default=8)
    ap.add_argument("--cpu-views", type=int, default=2)
    ap.add_argument("--graph-replay", type=int, default=1, choices=[0, 1],
                    help="1: view_run replays a captured CUDA graph per view (library default); 0: launches every kernel")
    ap.add_argument("--precision", default="fast", choices=["exact", "fast"],
                    help="fast (default): the post chain spends the tolerance north_star grants the images (each kernel <= 1 fp16 ulp "
                         "from the oracle on the same inputs, 8-bit output <= 1 LSB), licensed by tests/test_precision_fast_gpu.py and "
                         "test_config4_64_frame_sequence_at_4k; exact: every image bit-identical to the CPU oracle.  Visibility and "
                         "light lists are bit-exact in both")
    ap.add_argument("--forward", action="store_true",
                    help="widened path (SURVEY 8.6 row f-1): depth and HDR colour are produced on the device from the cull objects "
                         "and the tile light lists instead of being uploaded; not the headline workload")
    ap.add_argument("--two-phase", action="store_true",
                    help="with --forward (row f-3): two-phase occlusion culling around the device rasteriser, phase 1 against the "
                         "previous step's Hi-Z pyramid of the same view slot")
    ap.add_argument("--velocity", action="store_true",
                    help="supply a per-pixel RG16F velocity buffer to the TAA (the headline workload has none, PIN-12)")
    ap.add_argument("--instancing", action="store_true",
                    help="widened path (SURVEY 8.6 row f-2): also run LOD selection + instance batching after the cull; "
                         "not part of the headline workload, reported under \"widened\"")
    args = ap.parse_args()
    if args.two_phase and not args.forward:
        ap.error("--two-phase needs --forward")
    if args.impl == "reference":
        run_reference(args)
    else:
        run_cuda(args)


if __name__ == "__main__":
    main()
