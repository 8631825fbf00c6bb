#!/usr/bin/env python
"""bench.py — throughput of the trace loop (BASELINE.json metric: Mrays/s & spp/s at 1080p, 2 bounces).

    python bench.py --gpus N --steps K --warmup W        (N>1: launched under torch.distributed.run)
    python bench.py --impl reference ...                 (CPU oracle on the host cores)

A step = one frame of the hot path (raygen -> [trace -> shade]* -> image) of the synthetic
Sponza-scale scene (config C2: ~260 k triangles, 25 textured PBR materials, 1920x1080, 1 spp,
2 ray segments) with frame_count = step index.  At N>1 every rank renders its own frame_count
(sample-range partitioning: rank r renders frames r, r+N, ...) into its own accumulation buffer and
the per-rank accumulation buffers are combined with ONE NCCL reduce(sum) to rank 0 — the library's
pt_reduce_accum — inside the timed region, after the last step (weak scaling).  The `strong` object of
the line is config C3 (3840x2160, 64 spp, 8 segments) with every frame tile-partitioned over the N
ranks and reduced per frame (pt_reduce_image).  torch.distributed is used for the launch plumbing only
(barrier, max-over-ranks of the timings, shipping the 128-byte NCCL id).
Prints ONE JSON line on rank 0.
"""
import argparse
import ctypes as C
import json
import math
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "vulkan-pathtracer_b200"))

from tools import scene as S  # noqa: E402

WORKLOADS = {
    # name: (builder kwargs, W, H, spp, segments)
    "c1": ("cornell", 512, 512, 1, 2),
    "c2": ("sponza_scale", 1920, 1080, 1, 2),
    "c3": ("sponza_scale", 3840, 2160, 64, 8),
    "c4": ("instanced", 1920, 1080, 4, 4),
    "c5": ("heavy_material", 1920, 1080, 16, 6),
}
WORKLOAD_DESC = {
    "c1": "C1 synthetic Cornell box 512x512 1spp 2 segments",
    "c2": "C2 synthetic Sponza-scale glTF (~260k tris, 25 textured PBR materials) 1920x1080 1spp 2 segments",
    "c3": "C3 synthetic Sponza-scale 3840x2160 64spp 8 segments",
    "c4": "C4 1000 instances x 10k-tri mesh 1920x1080 4spp 4 segments",
    "c5": "C5 ~1M tris, 4096^2 textures, 1920x1080 16spp 6 segments",
}


def build_scene(name):
    t = time.time()
    sb = getattr(S, WORKLOADS[name][0])(1234)
    ps = sb.pack()
    return ps, time.time() - t


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            return json.load(f), "measured"
    return {"hbm_gbs": 6650.0}, "fallback"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons.  The process is started before the warm-up (it needs ~0.2 s to print its first
    row) and every row is time-stamped on arrival; stop(t0, t1) keeps the rows that fall inside the timed region, or — when
    the region is shorter than the 20 ms sampling period allows — the rows nearest to it, and says which."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.rows = []
        self.proc = None
        self.gpu_index = gpu_index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "20",
                                          "-i", str(self.gpu_index)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), line.strip()))

    def stop(self, t0=None, t1=None):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.05)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        rows, where = self.rows, "whole run"
        if t0 is not None:
            inside = [r for r in self.rows if t0 <= r[0] <= t1 + 0.03]
            if inside:
                rows, where = inside, "timed region"
            else:
                near = sorted(self.rows, key=lambda r: min(abs(r[0] - t0), abs(r[0] - t1)))[:3]
                rows, where = near, "nearest to the timed region (region shorter than the sampling period)"
        sm, mx, reasons = [], [], set()
        for _, r in rows:
            f = [x.strip() for x in r.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm), "window": where}


def cpu_arm(ps, name, seconds=None, steps=None, warmup=0):
    """The CPU arm of the comparison on all host threads, on a bounded sample of the workload: every `share`-th interleaved
    32x8 tile of full-size frames at 1 spp (Mrays/s is a rate).  kind "reference": the reference's OWN shader sources
    (ray_gen.rgen, closest_hit.rchit, miss.rmiss, triangle/material/brdf/rand.glsl) compiled as C++ into oracle/_ref
    (oracle/ref_glsl/), one raygen invocation per pixel, with the oracle's BVH2 closest-hit query standing in for the part
    the Vulkan driver owns (acceleration-structure traversal has no source).  kind "port": the oracle's restatement, when
    oracle/_ref was never built.  Returns (Mrays/s, dict for the JSON line, ms per full frame)."""
    from oracle import orc
    _, W, H, spp, seg = WORKLOADS[name]
    osc = orc.OracleScene(ps, True)
    nt = orc.lib().orc_hardware_threads()
    ref = orc.ref_lib()
    use_ref = ref is not None and hasattr(ref, "ref_render")
    share = max(1, int(round((W * H * seg / 2.0e6) / max(1, nt) / 3.0)))
    rays_total, t_total, frames = 0, 0.0, 0
    i = 0
    while True:
        cfg = S.render_config(W, H, 1, seg, rank=0, world=share, partition=1 if share > 1 else 0)
        cam = S.camera_default(W, H, i)
        t0 = time.perf_counter()
        if use_ref:
            _, rays = osc.ref_render(cfg, cam, True, nt)
        else:
            _, _, rays = osc.render(cfg, cam, True, nt, ids=False)
        dt = time.perf_counter() - t0
        if i >= warmup:
            rays_total += rays; t_total += dt; frames += 1
        i += 1
        if (steps is not None and frames >= steps) or (seconds is not None and t_total >= seconds) or frames >= 4096:
            break
    mrays = rays_total / t_total / 1e6
    sample = (f"{frames} frames x 1/{share} of the interleaved 32x8 tiles of {WORKLOAD_DESC[name]} at 1 spp, {nt} threads, {t_total:.2f} s; "
              + ("the reference's own GLSL (ray_gen.rgen + includes) compiled as C++ (oracle/_ref), oracle BVH2 closest-hit behind traceRayEXT"
                 if use_ref else "CPU oracle (C++ restatement, binned-SAH BVH2)"))
    info = {"value": mrays, "unit": "Mrays/s", "cores": nt, "kind": "reference" if use_ref else "port", "sample": sample}
    return mrays, info, 1e3 * t_total / frames * share


def run_reference(args, rank, world):
    """--impl reference: the reference's Vulkan pipeline cannot run here (no Zig, no Vulkan: SURVEY.md 8c); its shader code
    can — see cpu_arm()."""
    if rank != 0:
        return 0
    name = args.workload
    ps, _ = build_scene(name)
    _, W, H, spp, seg = WORKLOADS[name]
    mrays, info, ms_frame = cpu_arm(ps, name, steps=args.steps, warmup=args.warmup)
    line = {"impl": "reference", "metric": "Mrays/s", "value": mrays, "unit": "Mrays/s", "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_frame, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": WORKLOAD_DESC[name], "note": "the reference's Vulkan/lavapipe path is unavailable on this box; its shader sources compiled as C++ are timed on the host cores instead (see cpu_baseline.sample); ms_per_step extrapolated to the full frame"},
            "cpu_baseline": info,
            "e2e": {"value": mrays, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "spp_per_s": 1e3 / ms_frame}
    print(json.dumps(line))
    return 0


class DevArray:
    """__cuda_array_interface__ view of a device buffer owned by the renderer."""

    def __init__(self, ptr, shape):
        self.__cuda_array_interface__ = {"shape": shape, "typestr": "<f4", "data": (ptr, False), "version": 2}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="c2", choices=sorted(WORKLOADS))
    ap.add_argument("--flags", type=int, default=0)
    ap.add_argument("--partition", default="frames", choices=["frames", "tiles", "tiles-peer"],
                    help="N>1: frames = rank r renders frame_count r, r+N, ... (weak scaling, default); tiles = every rank renders "
                         "its interleaved 32x8 tiles of the SAME frame + NCCL reduce (strong scaling); tiles-peer = same, but owned "
                         "pixels are stored straight into rank 0's image over NVLink (CUDA IPC), one 4-byte NCCL all-reduce as fence")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-strong", action="store_true", help="skip the strong-scaling leg (C3, tiles + reduce per frame)")
    ap.add_argument("--strong-steps", type=int, default=3)
    ap.add_argument("--no-l2-flush", action="store_true")
    ap.add_argument("--single-frame-in-flight", action="store_true", help="e2e leg without the sibling frame contexts")
    ap.add_argument("--frames-in-flight", type=int, default=2, help="e2e leg: frame contexts over one device scene (pt_create_sibling), one per "
                    "swapchain image in the reference (main.zig:209-217)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 0)
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        return run_reference(args, rank, world)

    # stdout carries exactly ONE JSON line: libraries that print there (NCCL's version banner) go to stderr
    sys.stdout.flush()
    json_fd = os.dup(1)
    os.dup2(2, 1)

    import torch
    import torch.distributed as dist
    import b200pt

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (the b200 arm has no CPU fallback)")
    torch.cuda.set_device(local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    name = args.workload
    _, W, H, spp, seg = WORKLOADS[name]
    sampler = ClockSampler(local_rank)      # started now: nvidia-smi needs a few hundred ms before its first row, the timed region
    sampler.start()                         # of a 20-step run lasts 30 ms
    ps, t_gen = build_scene(name)
    t0 = time.time()
    tiles = world > 1 and args.partition != "frames"
    peer = world > 1 and args.partition == "tiles-peer"
    r = b200pt.Renderer(ps, W, H, spp, seg, device=local_rank, flags=args.flags, rank=rank if tiles else 0, world=world if tiles else 1,
                        partition=b200pt.PT_PARTITION_TILES if tiles else b200pt.PT_PARTITION_NONE)
    t_create = time.time() - t0
    if peer:
        # rank 0 exports its image buffer; the others map it and their K-shade writes owned pixels into it
        handle = torch.zeros(64, dtype=torch.uint8, device="cuda")
        if rank == 0:
            hb = (C.c_uint8 * 64)()
            b200pt.check(b200pt.lib().pt_ipc_export_image(r.handle, hb))
            handle.copy_(torch.tensor(list(hb), dtype=torch.uint8))
        dist.broadcast(handle, src=0)
        if rank != 0:
            hb = (C.c_uint8 * 64)(*handle.cpu().tolist())
            b200pt.check(b200pt.lib().pt_ipc_open_root_image(r.handle, hb))
        fence = torch.zeros(1, dtype=torch.int32, device="cuda")
    def comm_bootstrap(renderer):
        """pt_comm_init on every rank: rank 0 makes the NCCL id, torch.distributed ships its 128 bytes."""
        idt = torch.zeros(128, dtype=torch.uint8, device="cuda")
        if rank == 0:
            idt.copy_(torch.tensor(list(b200pt.comm_unique_id()), dtype=torch.uint8))
        dist.broadcast(idt, src=0)
        renderer.comm_init(rank, world, bytes(idt.cpu().tolist()))

    if world > 1 and not peer:
        comm_bootstrap(r)
    frames_mode = world > 1 and not tiles
    if frames_mode:
        r.set_accumulate(True)      # every rank sums its own frames; one reduce of the accumulation buffers at the end
    st0 = r.stats()
    # a non-default torch stream: its handle is passed to pt_render so that torch.cuda.Event timing and
    # NCCL ordering see the renderer's kernels (the NULL handle would mean "the renderer's own stream")
    stream = torch.cuda.Stream()
    torch.cuda.set_stream(stream)
    sp = C.c_void_p(stream.cuda_stream)
    assert sp.value, "expected a non-default CUDA stream"
    img_t = torch.as_tensor(DevArray(r.device_image_ptr(), (H, W, 4)), device=f"cuda:{local_rank}")
    flush = None if args.no_l2_flush else torch.empty(256 * 1024 * 1024 // 4, dtype=torch.float32, device="cuda")
    total_steps = args.warmup + args.steps

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def frame_of(i):
        return i if tiles else i * world + rank

    def combine():
        if world == 1 or frames_mode:
            return
        if peer:
            dist.all_reduce(fence)      # stream-ordered fence: every rank's kernels (and their peer stores) precede it
        else:
            r.reduce_image(0, sp)       # ncclReduce(sum) to rank 0 inside libb200pt, on the timed stream

    def render_frame(cam):
        r.record(cam, sp)
        combine()

    def step(i):
        render_frame(S.camera_default(W, H, frame_of(i)))

    # ---------------- device-timed region: inputs resident, per-step CUDA events, L2 flushed between steps
    r.set_profiling(True)
    for i in range(args.warmup):
        step(i)
    if frames_mode:
        r.reduce_accum(0, sp)           # warm-up of the collective too
        r.reset_accum()
    barrier()
    t_region0 = time.time()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    trace_ms = shade_ms = other_ms = 0.0
    launches = 0
    rays_local = 0
    for k in range(args.steps):
        if flush is not None:
            flush.fill_(float(k))
        ev[k][0].record(stream)
        step(args.warmup + k)
        ev[k][1].record(stream)
        # stats of this frame (synchronises; outside the event bracket)
        s = r.stats()
        trace_ms += s.ms_trace; shade_ms += s.ms_shade; other_ms += s.ms_other
        launches += s.kernel_launches
        rays_local += s.rays_traced
    reduce_ms = 0.0
    if frames_mode:
        # the combine of the sample-range split: per-rank accumulation buffers -> rank 0, once, inside the timed region
        ra, rb = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ra.record(stream)
        r.reduce_accum(0, sp)
        rb.record(stream)
        launches += 1
    barrier()
    t_region1 = time.time()
    clocks = sampler.stop(t_region0, t_region1)
    if frames_mode:
        reduce_ms = ra.elapsed_time(rb)
    dev_ms = sum(a.elapsed_time(b) for a, b in ev) + reduce_ms
    t = torch.tensor([dev_ms, float(rays_local)], dtype=torch.float64, device="cuda")
    if world > 1:
        tmax = t.clone(); dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        tsum = t.clone(); dist.all_reduce(tsum, op=dist.ReduceOp.SUM)
        dev_ms, rays_total = float(tmax[0]), float(tsum[1])
    else:
        rays_total = float(rays_local)
    r.set_profiling(False)
    mrays = rays_total / (dev_ms * 1e-3) / 1e6
    frames_done = args.steps * (1 if tiles else world)
    spp_s = spp * frames_done / (dev_ms * 1e-3)

    # ---------------- end-to-end: public API with host buffers (camera in, image out through pinned memory)
    host_img = torch.empty((H, W, 4), dtype=torch.float32, pin_memory=True)
    # One GPU: two frames in flight, like the reference's one command buffer + fence per swapchain image
    # (main.zig:209-217,255-259): a sibling context shares the device scene and has its own stream, path state,
    # image and pinned destination, so the tail of frame k overlaps ray generation + primary traversal of frame k+1.
    # N > 1, frame shares: the same on every rank (no per-step collective), both contexts accumulate; after the last step the
    # sibling's accumulation is folded into the parent's, the parents' are reduced to rank 0 and rank 0 reads the mean.
    in_flight = max(2, args.frames_in_flight) if ((world == 1 or frames_mode) and not args.single_frame_in_flight) else 1
    ctx, hosts = [r], [host_img]
    for _ in range(in_flight - 1):
        c = r.sibling()
        if frames_mode:
            c.set_accumulate(True)
        h = torch.empty((H, W, 4), dtype=torch.float32, pin_memory=True)
        for k in range(2):      # untimed: the sibling creates its copy stream and events on first use
            c.record(S.camera_default(W, H, k))
            b200pt.check(b200pt.lib().pt_read_rgba32f_overlapped(c.handle, h.data_ptr()))
        c.sync()
        ctx.append(c); hosts.append(h)
    if frames_mode:
        for c in ctx:
            c.reset_accum()
    barrier()
    t0 = time.perf_counter()
    e2e_rays = 0
    for k in range(args.steps if in_flight >= 2 else 0):
        cam = S.camera_default(W, H, frame_of(args.warmup + k))     # host camera struct (68 B) crosses in pt_render
        c = ctx[k % in_flight]
        c.record(cam)
        if rank == 0:       # the step's result crosses to the host on rank 0 (33.2 MB per step, as d2h_bytes_per_step says)
            b200pt.check(b200pt.lib().pt_read_rgba32f_overlapped(c.handle, hosts[k % in_flight].data_ptr()))
    if in_flight >= 2:
        for c in ctx:
            c.sync()
    for k in range(args.steps if (in_flight == 1 and not frames_mode) else 0):
        cam = S.camera_default(W, H, frame_of(args.warmup + k))     # host camera struct (68 B) crosses in pt_render
        render_frame(cam)
        if rank == 0:
            if peer:    # the fence below is ordered on `sp`, so the copy has to be too
                b200pt.check(b200pt.lib().pt_read_rgba32f_async(r.handle, host_img.data_ptr(), sp))
            else:       # copy of frame k overlaps ray generation + primary traversal of frame k+1
                b200pt.check(b200pt.lib().pt_read_rgba32f_overlapped(r.handle, host_img.data_ptr()))
        if peer:
            dist.all_reduce(fence)      # the next frame's peer stores must not overtake rank 0's read-back
    for k in range(args.steps if (in_flight == 1 and frames_mode) else 0):
        r.record(S.camera_default(W, H, frame_of(args.warmup + k)))
        if rank == 0:
            b200pt.check(b200pt.lib().pt_read_rgba32f_overlapped(r.handle, host_img.data_ptr()))
    if frames_mode:
        for c in ctx[1:]:
            r.add_accum(c)
        r.reduce_accum(0)               # on the library's communication stream, behind this rank's last frame
        if rank == 0:
            b200pt.check(b200pt.lib().pt_read_accum_mean(r.handle, host_img.data_ptr(), args.steps * world))   # synchronises
    r.sync()            # the last pipelined read-back
    barrier()
    e2e_s = time.perf_counter() - t0
    te = torch.tensor([e2e_s], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
    e2e_s = float(te[0])
    e2e_mrays = rays_total / e2e_s / 1e6   # same frames => same ray counts as the device-timed region

    # ---------------- roofline of the dominant kernel (K-trace): algorithmic bytes / event-timed duration
    pk, pk_src = peaks()
    roof = cpu = None
    if rank == 0:
        nodes, tris, crays = r.count_traversal()            # independent one-thread-per-ray walker, last frame
        bytes_per_ray = (80.0 * nodes + 48.0 * tris) / max(1, crays)
        n_trace_launches = args.steps * seg if spp == 1 else None
        achieved = bytes_per_ray * rays_local / (trace_ms * 1e-3) / 1e9 if trace_ms > 0 else 0.0
        # DRAM bytes per K-trace launch from the committed `ncu --set full` capture of this workload (None if absent)
        traffic = warp_inst = lanes = None
        try:
            with open(os.path.join(ROOT, "profiles", "ncu_traffic.json")) as f:
                tj = json.load(f)
            if tj.get("workload") == args.workload and tj.get("flags", 0) == args.flags:
                kt = tj["k_trace_persistent"]
                traffic = kt["dram_bytes_per_launch"]
                warp_inst = kt.get("warp_inst_per_frame")           # smsp__inst_executed.sum over the frame's K-trace launches
                lanes = kt.get("lanes_per_inst")                    # smsp__thread_inst_executed_per_inst_executed, instruction-weighted
        except (OSError, KeyError, ValueError):
            pass
        # the bounds K-trace actually sits under (SURVEY 8(d) "which roofline"): the traversal working set is L2-resident, so
        # the memory bound is the L2 read bandwidth, measured here and now (32 MiB buffer read 100 times through __ldcg);
        # and the kernel is instruction-issue-bound, so the issue rate (4 warp instructions per SM per clock) is stated too
        l2_gbs, n_sms = b200pt.read_bandwidth(32 << 20, 100, local_rank)
        sm_mhz = clocks.get("sm_mhz") or pk.get("sm_max_mhz", 1965.0)
        issue_peak = n_sms * 4 * sm_mhz * 1e6                       # warp instructions per second
        issue_frac = (warp_inst * args.steps / (trace_ms * 1e-3)) / issue_peak if (warp_inst and trace_ms > 0) else None
        # bloat guard: one root-to-leaf descent of a perfect 8-ary tree with 3 triangles per leaf
        per_mesh = [sum(ps.primitives[p].triangle_count for p in range(m.start, m.end)) for m in ps.meshes]
        t_blas, n_inst = max(per_mesh + [1]), max(1, len(ps.instances))
        ceil_log8 = lambda v: 0 if v <= 1 else int(math.ceil(math.log(v, 8) - 1e-9))
        a_floor = 80.0 * (ceil_log8(t_blas / 3.0) + ceil_log8(n_inst)) + 48.0 * 3
        roof = {"bound": "hbm", "achieved": achieved, "peak": pk["hbm_gbs"], "unit": "GB/s", "frac": achieved / pk["hbm_gbs"],
                "traffic": traffic, "kernel": "k_trace_persistent", "peak_source": pk_src,
                "algorithmic_bytes_per_ray": bytes_per_ray, "a_floor_bytes_per_ray": a_floor, "a_trav_over_floor": bytes_per_ray / a_floor, "nodes_per_ray": nodes / max(1, crays), "tris_per_ray": tris / max(1, crays),
                "trace_ms_per_step": trace_ms / args.steps, "shade_ms_per_step": shade_ms / args.steps, "other_ms_per_step": other_ms / args.steps,
                "trace_mrays_per_s": rays_local / (trace_ms * 1e-3) / 1e6 if trace_ms > 0 else 0.0,
                "l2_peak_gbs": l2_gbs, "l2_frac": achieved / l2_gbs if l2_gbs > 0 else None, "l2_peak_source": "measured in this run (pt_debug_read_bandwidth, 32 MiB x 100 passes)",
                "issue_frac": issue_frac, "lanes_per_inst": lanes, "useful_issue_frac": issue_frac * lanes / 32.0 if (issue_frac and lanes) else None,
                "issue_peak_gwarp_inst_s": issue_peak * 1e-9, "issue_source": "warp instructions per frame from the committed ncu capture (profiles/ncu_traffic.json) / K-trace time of this run",
                "note": "BVH (nodes+triangles) is L2-resident on B200: frac is against the HBM copy peak (the contract's denominator), l2_frac against the measured L2 read "
                        "bandwidth, issue_frac against the issue rate -- the binding one (see DESIGN.md section 5)"}
        if not args.no_cpu_baseline and world == 1:
            _, cpu, _ = cpu_arm(ps, name, seconds=12.0)
    # ---------------- strong scaling (north_star: ">= 7x at 8 GPUs", BASELINE config 3): C3 = the same scene at 3840x2160,
    # 64 spp, 8 segments; every frame is split into interleaved 32x8 tiles over the N ranks and reduced to rank 0 per frame
    # (pt_reduce_image) inside the timed region.  Total work is fixed: value(N) / value(1) is the strong-scaling curve.
    strong = None
    if name == "c2" and not args.no_strong and args.flags == 0:
        _, W3, H3, spp3, seg3 = WORKLOADS["c3"]
        r3 = b200pt.Renderer(ps, W3, H3, spp3, seg3, device=local_rank, rank=rank, world=world,
                             partition=b200pt.PT_PARTITION_TILES if world > 1 else b200pt.PT_PARTITION_NONE)
        if world > 1:
            comm_bootstrap(r3)
        ev3 = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.strong_steps)]
        rays3 = 0
        for i in range(1 + args.strong_steps):
            if i == 1:
                barrier()
            k = i - 1
            if k >= 0:
                ev3[k][0].record(stream)
            r3.record(S.camera_default(W3, H3, i), sp)
            if world > 1:
                r3.reduce_image(0, sp)
            if k >= 0:
                ev3[k][1].record(stream)
                rays3 += r3.stats().rays_traced          # synchronises, outside the bracket
        barrier()
        t3 = torch.tensor([sum(a.elapsed_time(b) for a, b in ev3), float(rays3)], dtype=torch.float64, device="cuda")
        if world > 1:
            t3m = t3.clone(); dist.all_reduce(t3m, op=dist.ReduceOp.MAX)
            t3s = t3.clone(); dist.all_reduce(t3s, op=dist.ReduceOp.SUM)
            ms3, rays3 = float(t3m[0]), float(t3s[1])
        else:
            ms3 = float(t3[0])
        strong = {"workload": WORKLOAD_DESC["c3"], "partition": "none" if world == 1 else "interleaved 32x8 tiles of every frame + NCCL reduce(sum) to rank 0 per frame (pt_reduce_image)",
                  "scaling": "strong", "n_gpus": world, "steps": args.strong_steps, "warmup": 1, "ms_per_step": ms3 / args.strong_steps,
                  "value": rays3 / (ms3 * 1e-3) / 1e6, "unit": "Mrays/s", "rays_per_step": rays3 / args.strong_steps,
                  "spp_per_s": spp3 * args.strong_steps / (ms3 * 1e-3)}
        r3.deinit()
    if rank == 0:
        line = {"metric": "Mrays/s", "value": mrays, "unit": "Mrays/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": dev_ms / args.steps, "higher_is_better": True, "scaling": "strong" if tiles else "weak", "vs_baseline": None, "dtype": "f32",
                "data": "synthetic",
                "config": {"workload": WORKLOAD_DESC[name], "partition": ("none" if (world == 1 and not tiles) else "frames (rank r renders frame_count r, r+N, ... into its own accumulation buffer) + one NCCL reduce(sum) of the accumulation buffers to rank 0 after the last step (pt_reduce_accum), inside the timed region" if not tiles
                                         else "interleaved 32x8 tiles of one frame, peer stores into rank 0 over NVLink + 4-byte NCCL fence" if peer
                                         else "interleaved 32x8 tiles of one frame + NCCL reduce(sum) to rank 0"),
                           "l2": "flushed between timed steps (256 MiB write)" if flush is not None else "not flushed; per-step path state (365 MB) exceeds L2",
                           "triangles": int(st0.num_triangles), "bvh8_nodes": int(st0.bvh_nodes), "bvh_build_ms": st0.ms_build,
                           "scene_gen_s": t_gen, "create_s": t_create},
                "spp_per_s": spp_s, "rays_per_step": rays_total / args.steps / (1 if tiles else world),
                "e2e": {"value": e2e_mrays, "unit": "Mrays/s", "h2d_bytes_per_step": C.sizeof(S.PtCamera), "d2h_bytes_per_step": W * H * 16,
                        "ms_per_step": 1e3 * e2e_s / args.steps, "spp_per_s": spp * frames_done / e2e_s, "frames_in_flight": in_flight},
                "gpu_launches": int(launches), "clocks": clocks, "roofline": roof, "cpu_baseline": cpu, "strong": strong}
        sys.stdout.flush()
        os.write(json_fd, (json.dumps(line) + "\n").encode())
    for c in ctx[1:]:
        c.deinit()          # siblings go first: the scene belongs to r
    r.deinit()
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
