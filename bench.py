#!/usr/bin/env python
"""
bench.py — BASELINE.json's metric on BASELINE.json's config.

  metric   srn_samples_per_sec: useful SRN network samples per second while rendering the paper-default fV-SRN
           (hidden 32x4, 14 Fourier features, 32^3 x 16 fp16 latent grid, ReLU) as a 1024^2 DVR frame, step 1/512,
           piecewise TF, Beer-Lambert, early termination ("C3", BASELINE.json configs[2] — the config the metric's
           "ms/frame (1024^2 DVR)" is quoted on). A "useful" sample is one the marcher needs (inside the box, ray not yet
           saturated); lock-step padding samples are reported separately and never counted. ms/frame is `ms_per_step`
           (one view per GPU per step).
  value    device-timed, inputs resident in HBM.   e2e: the same frame through fvsrn_render_dvr_host (host camera + TF
           in, host image out, copies inside the timed region).
  N GPUs   weak scaling (default): every rank renders its own view of a camera sweep per step (view-sharded multi-view
           batch); the finished pixels (RGBA8 written by the DVR kernel itself) are collected on rank 0 over NCCL inside the
           timed region, the transfer of step k overlapping the render of step k+1 (two buffers).
           --sharding tiles: strong scaling - ONE frame per step, split into contiguous row bands of equal estimated
           march cost (BASELINE C4: 2048^2, 4 spp); bands are received in place on rank 0.
  --impl reference   the reference's CPU PyTorch path (oracle/torch_ref.py, pinned to the reference by tests/golden)
           on all host cores, on a bounded centre crop of the same frame.
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for _p in (ROOT, os.path.join(ROOT, "oracle")):
    if _p not in sys.path:
        sys.path.insert(0, _p)

import numpy as np  # noqa: E402

WORKLOADS = {
    # name: network kwargs, image size, stepsize, tensor FLOP per sample
    "C3": dict(net=dict(hidden=32, num_layers=4, grid_res=32, activation="ReLU", seed=42, last_gain=6.0),
               size=1024, stepsize=1.0 / 512, flop=7168, desc="fV-SRN 32x4, 14 Fourier, 32^3x16 fp16 grid, ReLU; DVR 1024^2, step 1/512, "
               "piecewise TF (plume100), Beer-Lambert, early termination"),
    "C3-snake": dict(net=dict(hidden=32, num_layers=4, grid_res=32, activation="SnakeAlt:2", seed=42, last_gain=12.0),
                     size=1024, stepsize=1.0 / 512, flop=7168, desc="as C3 with SnakeAlt:2 activations"),
    "C1": dict(net=dict(hidden=32, num_layers=4, grid_res=16, activation="ReLU", seed=42, last_gain=6.0),
               size=256, stepsize=0.005, flop=7168, desc="fV-SRN 32x4, 16^3x16 grid; DVR 256^2, step 0.005"),
    "C4": dict(net=dict(hidden=64, num_layers=6, grid_res=64, activation="ReLU", seed=42, hidden_gain=2.0),
               size=2048, stepsize=1.0 / 512, flop=43008, spp=4, desc="fV-SRN 64x6, 30 Fourier, 64^3x16 grid; DVR 2048^2, 4 spp"),
    "C4-1spp": dict(net=dict(hidden=64, num_layers=6, grid_res=64, activation="ReLU", seed=42, hidden_gain=2.0),
                    size=2048, stepsize=1.0 / 512, flop=43008, desc="fV-SRN 64x6, 30 Fourier, 64^3x16 grid; DVR 2048^2, 1 spp"),
    "C5": dict(net=dict(hidden=32, num_layers=4, grid_res=32, activation="ReLU", seed=42, last_gain=6.0, time_keyframes=8),
               size=512, stepsize=1.0 / 512, flop=7168, views=8, desc="time-dependent fV-SRN 32x4 (8 keyframe grids 32^3x16, temporally "
               "interpolated per view), 512^2 camera sweep, 8 views per GPU per step"),
}


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return dict(tflops=float(d["bf16_tflops"]), tflops_sustained=float(d.get("bf16_tflops_sustained", d["bf16_tflops"])),
                    hbm=float(d["hbm_gbs"]), source="measured (MEASURED_PEAKS.json)")
    return dict(tflops=1590.0, tflops_sustained=1400.0, hbm=6650.0, source="fallback (B200_PROFILING.md)")


class ClockSampler:
    """nvidia-smi clocks + throttle reasons sampled DURING the timed region (B200_PROFILING.md recipe)."""
    Q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        self.rows, self._stop, self.index = [], threading.Event(), index
        self._t = threading.Thread(target=self._run, daemon=True)

    def _run(self):
        while not self._stop.is_set():
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits"],
                                     capture_output=True, text=True, timeout=5).stdout.strip()
                if out:
                    self.rows.append([c.strip() for c in out.split(",")])
            except Exception:
                pass
            self._stop.wait(0.1)

    def start(self): self._t.start()

    def stop(self):
        self._stop.set(); self._t.join(timeout=6)
        sm = sorted(int(r[0]) for r in self.rows if r and r[0].isdigit())
        mx = max([int(r[1]) for r in self.rows if len(r) > 1 and r[1].isdigit()] or [0])
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(len(r) > 2 + i and r[2 + i].lower().startswith("active") for r in self.rows)]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx or None, "reasons": reasons, "samples": len(sm)}


def make_scene(workload):
    import fvsrn_b200  # noqa: F401
    from fvsrn_b200 import scenes
    w = WORKLOADS[workload]
    net = scenes.make_network(**w["net"])
    tf = scenes.piecewise_tf_tensor()
    return net, tf, scenes


def crop_for_cpu(size, px):
    """Centre crop of the frame used as the bounded CPU sample."""
    x0 = (size - px) // 2
    return (x0, x0, px, px)


# ------------------------------------------------------------------------------------------------ reference arm (CPU)
def run_reference(args):
    """The reference's own CPU path: PyTorch network + PyTorch marcher (restated in oracle/torch_ref.py, pinned to the
    reference by tests/golden), all host threads, on a bounded centre crop of the workload's frame."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    import torch
    import torch_ref
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    w = WORKLOADS[args.workload]
    net, tf, scenes = make_scene(args.workload)
    cam = scenes.camera_reference_frame(**scenes.DEFAULT_CAMERA)
    tn = torch_ref.TorchNetwork(net)
    kw = dict(width=w["size"], height=w["size"], camera=cam, fov_y=scenes.DEFAULT_FOV_Y, stepsize=w["stepsize"], tf=tf)
    rate = 1.0
    for _ in range(max(1, args.warmup)):
        t0 = time.perf_counter()
        _, ex, _ = torch_ref.render_dvr(tn, **dict(kw, tile=crop_for_cpu(w["size"], 16)))
        rate = ex / (time.perf_counter() - t0)
    px = adaptive_crop(args, w, rate)
    kw["tile"] = crop_for_cpu(w["size"], px)
    t0 = time.perf_counter(); useful = 0; executed = 0
    for _ in range(args.steps):
        _, ex, us = torch_ref.render_dvr(tn, **kw)
        useful += us; executed += ex
    dt = time.perf_counter() - t0
    value = useful / dt
    sample = f"centre {px}x{px} crop of the {w['size']}^2 frame per step (torch marcher executes {executed // args.steps} samples, " \
             f"{useful // args.steps} useful)"
    line = {"impl": "reference", "metric": "srn_samples_per_sec", "value": value, "unit": "samples/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"{args.workload}: {w['desc']}", "sample": sample},
            "cpu_baseline": {"value": value, "unit": "samples/s", "cores": cores, "kind": "port", "sample": sample},
            "e2e": {"value": value, "unit": "samples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    _emit(line)
    return 0


def adaptive_crop(args, w, rate_executed):
    """Crop edge so that the CPU leg stays bounded (~`cpu_budget_s` in total)."""
    max_steps = int(1.75 / w["stepsize"])
    px = int((args.cpu_budget_s * rate_executed / max(1, args.steps) / max_steps) ** 0.5)
    return max(16, min(args.cpu_crop, (px // 8) * 8))


# ------------------------------------------------------------------------------------------------ B200 arm
FMT_BYTES = {"rgba8": 4, "rgba_f16": 8, "rgba_f32": 16}


def _alloc_image(torch, fmt, n, h, w):
    if fmt == "rgba8": return torch.empty((n, h, w), dtype=torch.int32, device="cuda")
    if fmt == "rgba_f16": return torch.empty((n, h, w, 4), dtype=torch.float16, device="cuda")
    return torch.empty((n, 4, h, w), dtype=torch.float32, device="cuda")


def _image_shape(fmt, n, h, w):
    return {"rgba8": (n, h, w), "rgba_f16": (n, h, w, 4)}.get(fmt, (n, 4, h, w))


def _image_dtype(torch, fmt):
    return {"rgba8": torch.int32, "rgba_f16": torch.float16}.get(fmt, torch.float32)


def _image_strides(fmt, t, w):
    """out_stride_{b,c,y,x} of fvsrn_dvr_params in elements of the output format for a contiguous image tensor `t`
    (None: a full w x w frame that lives in another rank's memory)."""
    if t is None: return (4 * w * w, w * w, w, 1) if fmt == "rgba_f32" else (w * w, 0, w, 1)
    if fmt == "rgba_f32": return tuple(t.stride())
    per = 4 if fmt == "rgba_f16" else 1
    return (t.stride(0) // per, 0, w, 1)


def run_b200(args):
    import torch
    import torch.distributed as dist
    import fvsrn_b200  # noqa: F401
    from fvsrn_b200 import pyrenderer as pr, volnet, _lib, dist as fdist

    if not torch.cuda.is_available():
        raise RuntimeError("bench.py needs a CUDA device: the fV-SRN B200 path has no CPU fallback")
    world = int(os.environ.get("WORLD_SIZE", "1")); rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        raise RuntimeError(f"--gpus {args.gpus} but WORLD_SIZE={world}: launch with torch.distributed.run --nproc-per-node {args.gpus}")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    L = _lib.lib()
    w = WORKLOADS[args.workload]
    net, tf_np, scenes = make_scene(args.workload)
    S = w["size"]
    fmt = args.gather_format
    tiles = args.sharding == "tiles"
    peer_mode = world > 1 and args.collective == "peer"
    peer = None
    sn = pr.SceneNetwork.from_bytes(volnet.dumps(net))
    sn.kernel_mode = args.kernel_mode
    vol = pr.VolumeInterpolationNetwork(); vol.set_network(sn)
    ie = pr.ImageEvaluatorSimple(); ie.volume = vol
    V = 1 if tiles else int(w.get("views", 1))              # views per rank per step (view sharding) / one frame per step (tile sharding)
    cams = scenes.camera_sweep(1 if tiles else world * V)
    cam_np = np.ascontiguousarray(cams[0:1] if tiles else cams[rank * V:(rank + 1) * V])
    ie.camera.set_parameters(torch.from_numpy(cam_np))
    ie.samples_per_pixel = int(w.get("spp", 1))
    keep_alive = []
    T = int(w["net"].get("time_keyframes", 0))
    ie.ray_evaluator.tf.tensor.value = torch.from_numpy(tf_np).cuda()
    ie.ray_evaluator.stepsize = w["stepsize"]
    sn._ensure_prepared()
    counters = torch.zeros(2, dtype=torch.int64, device="cuda")
    cam_dev = ie.camera.get_parameters(); tf_dev = ie.ray_evaluator.tf._tensor()

    # ---- this rank's share of a step and where it lands. The DVR kernel writes the selected colour channel in the gather
    # format itself (fused colour extraction, fvsrn_dvr_params.output_format): no second pass, no 32 B/pixel tensor.
    if tiles:   # screen-band sharding of ONE frame: contiguous row bands of (nearly) equal estimated march cost
        bands = fdist.partition_rows_by_cost(scenes.estimate_row_cost(cam_np, scenes.DEFAULT_FOV_Y, S, S, net.box_min, net.box_size), world, align=8)
        if world > 1 and not args.no_rebalance:
            # feedback: the estimate knows the box, not the data (early termination, empty space). Render the estimated bands,
            # measure each rank's kernel time, rescale the estimate band by band and cut again - twice.
            row_cost = scenes.estimate_row_cost(cam_np, scenes.DEFAULT_FOV_Y, S, S, net.box_min, net.box_size)
            scratch = _alloc_image(torch, fmt, 1, S, S)
            for it in range(2):
                y0, bh = bands[rank]
                q = ie._params(S, S, cam_dev, tf_dev, _image_strides(fmt, scratch, S))
                q.output_format, q.channel_mode = pr.OUTPUT_FORMATS[fmt], 3
                q.tile_x0, q.tile_y0, q.tile_w, q.tile_h = 0, y0, S, bh
                ce = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
                ms = 0.0
                if bh > 0:
                    for k in range(3):
                        if k == 1: ce[0].record()
                        _lib.check(L.fvsrn_render_dvr(sn._h, C.byref(q), scratch.data_ptr(), C.c_void_p(torch.cuda.current_stream().cuda_stream)))
                    ce[1].record(); torch.cuda.synchronize()
                    ms = ce[0].elapsed_time(ce[1]) / 2
                all_ms = [None] * world
                dist.all_gather_object(all_ms, float(ms))
                bands, row_cost = fdist.rebalance_rows_by_time(row_cost, bands, all_ms, align=8)
            del scratch
        y0, bh = bands[rank]
        if peer_mode:
            peer = fdist.PeerFrames(_image_shape(fmt, 1, S, S), _image_dtype(torch, fmt), dst=0, slots=2)
            frames = [peer.tensor(slot) for slot in range(2)]                                          # None on ranks > 0: they hold no image
        else:
            frames = [_alloc_image(torch, fmt, 1, S if rank == 0 else max(bh, 1), S) for _ in range(2)]      # rank 0 holds the whole frame
        if fmt == "rgba_f32":
            raise RuntimeError("--sharding tiles gathers packed pixels: use --gather-format rgba8 or rgba_f16")
    else:
        bands = None
        if peer_mode:
            peer = fdist.PeerFrames(_image_shape(fmt, world * V, S, S), _image_dtype(torch, fmt), dst=0, slots=2)
            frames = [None, None]
            gathered = [peer.tensor(slot) for slot in range(2)]
        else:
            frames = [_alloc_image(torch, fmt, V, S, S) for _ in range(2)]
        gathered = gathered if peer_mode else [_alloc_image(torch, fmt, world * V, S, S) if (world > 1 and (rank == 0 or args.collective == "all_gather")) else None for _ in range(2)]
    params = []
    for slot in range(2):
        t = frames[slot]
        p = ie._params(S, S, cam_dev, tf_dev, _image_strides(fmt, t, S))
        p.output_format, p.channel_mode = pr.OUTPUT_FORMATS[fmt], 3
        p.counters_dev = counters.data_ptr()
        if tiles:
            p.tile_x0, p.tile_y0, p.tile_w, p.tile_h = 0, y0, S, bh
        if T > 1:   # per-view time: a golden-ratio sequence over the keyframe range, so that every rank's views need the temporal
            # interpolation of two keyframe grids (the expensive case) in the same proportion at every N
            times = [(T - 1) * (((rank * V + i) * 0.6180339887) % 1.0) for i in range(V)]
            vt = (C.c_float * V)(*times); keep_alive.append(vt)
            p.view_times_host = C.cast(vt, C.c_void_p)
        # band sharding: the kernel addresses pixels by their frame coordinates; a rank that only holds its band passes the
        # address its row y0 would have in a full frame
        if peer_mode:   # rank 0's frame, mapped into this process: band sharding writes rows [y0, y0 + bh) of frame 0, view sharding
            base = peer.address(slot, 0 if tiles else rank * V)      # writes views [rank V, (rank + 1) V)
        else:
            base = t.data_ptr() - ((y0 * S * FMT_BYTES[fmt]) if (tiles and rank != 0) else 0)
        params.append((p, base))
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")   # > 126 MB L2
    pending = [None, None]

    def wait(slot):
        if pending[slot] is not None:
            for wk in pending[slot]: wk.wait()
            pending[slot] = None

    def step(k, ev=None):
        """One step: render this rank's share (one fused kernel per view), then START the collection of the finished pixels
        on rank 0. Steps alternate between two buffers, so the transfer of step k overlaps the render of step k+1."""
        slot = k & 1
        wait(slot)
        if peer_mode: wait(slot ^ 1)     # rank 0 enqueues signal k-1 behind its reads of this slot's previous frame (PeerFrames protocol)
        p, base = params[slot]
        if ev: ev[0].record()
        if not tiles or bh > 0:
            _lib.check(L.fvsrn_render_dvr(sn._h, C.byref(p), base, stream))
        if ev: ev[1].record()
        if peer_mode:
            pending[slot] = [peer.signal(slot)]
        elif world > 1:
            if tiles:
                pending[slot] = fdist.gather_row_bands(frames[slot][:, :bh] if rank != 0 else None, bands, frames[slot] if rank == 0 else None, async_op=True)
            else:
                _, wk = fdist.gather_views(frames[slot], world * V, out=gathered[slot], collective=args.collective, async_op=True)
                pending[slot] = [wk]

    def barrier():
        wait(0); wait(1)
        if world > 1: dist.barrier()
        torch.cuda.synchronize()

    for k in range(max(3, args.warmup)):
        step(k); flush.zero_()
    barrier()
    counters.zero_()
    evs = [[torch.cuda.Event(enable_timing=True) for _ in range(2)] for _ in range(args.steps)]
    ev_all = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    sampler = ClockSampler(local) if rank == 0 else None
    if sampler: sampler.start()
    barrier()
    # the L2 flush between steps is hygiene, not work: its device time is measured on its own and subtracted
    fe = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    fe[0].record()
    for _ in range(args.steps): flush.zero_()
    fe[1].record(); torch.cuda.synchronize()
    flush_ms = fe[0].elapsed_time(fe[1])
    barrier()
    t_wall = time.perf_counter()
    ev_all[0].record()
    for k in range(args.steps):
        step(k, evs[k]); flush.zero_()
    wait(0); wait(1)
    ev_all[1].record()
    barrier()
    t_wall = time.perf_counter() - t_wall
    total_ms = max(1e-6, ev_all[0].elapsed_time(ev_all[1]) - flush_ms)
    kernel_ms = sum(e[0].elapsed_time(e[1]) for e in evs)
    cnt = counters.cpu().numpy().astype(np.float64)          # [useful, executed] over the K timed steps

    # ---- e2e: host buffers in, finished pixels on the host out, every step. N = 1: the reference-facing C-ABI call
    # (fvsrn_render_dvr_host: camera + TF from host memory, (B,4,H,W) fp32 colour - what render + extract_color return - into
    # pinned host memory). N > 1: per step every rank uploads its camera + TF, renders, the pieces are collected on rank 0
    # over NCCL and rank 0 copies the assembled frames to its host memory - N GPUs' frames on one host.
    host_cam = torch.from_numpy(cam_np.copy()).pin_memory(); host_tf = torch.from_numpy(np.ascontiguousarray(tf_np)).pin_memory()
    e2e_formats = {}
    if world == 1:
        def make_e2e(f):
            planes = 8 if f == "raw8_f32" else (4 if f == "rgba_f32" else 1)
            shape = (V, planes, S, S) if planes > 1 else ((V, S, S, 4) if f == "rgba_f16" else (V, S, S))
            dt = torch.float16 if f == "rgba_f16" else (torch.int32 if f == "rgba8" else torch.float32)
            host_img = torch.empty(shape, dtype=dt).pin_memory()
            q = ie._params(S, S, host_cam, host_tf, (planes * S * S, S * S, S, 1) if planes > 1 else (S * S, 0, S, 1))
            q.camera_matrix_dev = host_cam.data_ptr(); q.tf_dev = host_tf.data_ptr()
            q.output_format, q.channel_mode = pr.OUTPUT_FORMATS[f], 3
            q.view_times_host = params[0][0].view_times_host

            def run():
                _lib.check(L.fvsrn_render_dvr_host(sn._h, C.byref(q), host_img.data_ptr(), stream))
                return float(host_img.reshape(-1)[host_img.numel() // 2])        # read the result on the host
            return run, host_img.numel() * host_img.element_size()
        for f in ("rgba_f32", "raw8_f32", "rgba8"):
            run, nbytes = make_e2e(f)
            for _ in range(2): run()
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            for _ in range(args.steps): run(); flush.zero_()
            torch.cuda.synchronize()
            e2e_formats[f] = {"ms_per_step": max(1e-6, 1e3 * (time.perf_counter() - t0) - flush_ms) / args.steps, "d2h_bytes_per_step": nbytes}
        e2e_s = e2e_formats["rgba_f32"]["ms_per_step"] * 1e-3 * args.steps
        d2h_bytes = e2e_formats["rgba_f32"]["d2h_bytes_per_step"]
    else:
        res0 = (frames if tiles else gathered)
        host_img = torch.empty_like(res0[0], device="cpu").pin_memory() if rank == 0 else None
        d2h_bytes = host_img.numel() * host_img.element_size() if rank == 0 else 0

        def e2e_step(k):
            cam_dev.copy_(host_cam, non_blocking=True); tf_dev.copy_(host_tf, non_blocking=True)
            step(k)
            slot = k & 1
            wait(slot)
            if rank == 0:
                host_img.copy_(res0[slot], non_blocking=True)
                torch.cuda.synchronize()
                return float(host_img.reshape(-1)[host_img.numel() // 2])
            return 0.0
        for k in range(2): e2e_step(k)
        barrier()
        t0 = time.perf_counter()
        for k in range(args.steps): e2e_step(k); flush.zero_()
        barrier()
        e2e_s = max(1e-9, time.perf_counter() - t0 - flush_ms * 1e-3)
    clocks = sampler.stop() if sampler else None

    # ---- BASELINE configs[1] (C2): world-space batch query, 2^24 random positions through the same network
    # (IVolumeInterpolation::evaluate -> fvsrn_evaluate), device-resident positions; reported next to the DVR numbers
    batch_query = None
    if rank == 0 and not args.no_batch_query:
        n_q = 1 << 24
        gq = torch.Generator(device="cuda"); gq.manual_seed(7)
        pos = torch.rand((n_q, 3), generator=gq, device="cuda", dtype=torch.float32)     # 192 MiB > L2
        for _ in range(2): dens = vol.evaluate(pos)
        torch.cuda.synchronize()
        qe = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
        reps = 5
        qe[0].record()
        for _ in range(reps): dens = vol.evaluate(pos)
        qe[1].record(); torch.cuda.synchronize()
        q_ms = qe[0].elapsed_time(qe[1]) / reps
        batch_query = {"workload": "C2: 2^24 random positions, same network (fvsrn_evaluate)", "ms": q_ms, "value": n_q / (q_ms * 1e-3),
                       "unit": "samples/s", "tensor_frac_of_peak": n_q * w["flop"] / (q_ms * 1e-3) / 1e12 / peaks()["tflops"],
                       "checksum": float(dens.double().mean())}
        if world == 1 and not args.no_gpu_reference:
            # the reference's own batch-query kernel (EvaluateNoBatches + its tensor-core SRN, oracle/_ref PTX) on the same positions
            try:
                import ref_kernel
                if ref_kernel.available():
                    rk = ref_kernel.ReferenceEvalKernel(net, time=0.0)       # fvsrn_evaluate runs at the handle's current time (0)
                    ref_d, ref_ms = rk.evaluate(pos, timing_iters=3)
                    batch_query["reference_gpu_kernel"] = {"ms": ref_ms, "value": n_q / (ref_ms * 1e-3), "speedup_of_this_kernel": ref_ms / q_ms,
                                                           "max_abs_vs_reference": float((dens - ref_d).abs().max()), "gate": 1e-2}
                    del ref_d, rk
            except Exception as e:      # no build for this network shape (e.g. the 64x6 workloads)
                batch_query["reference_gpu_kernel"] = {"unavailable": str(e)[:200]}
        del pos, dens

    red = torch.tensor([total_ms, kernel_ms, e2e_s], dtype=torch.float64, device="cuda")
    tot = torch.tensor([cnt[0], cnt[1]], dtype=torch.float64, device="cuda")
    per_rank_kernel_ms = [kernel_ms / args.steps]
    if world > 1:
        allk = [torch.zeros(1, dtype=torch.float64, device="cuda") for _ in range(world)]
        dist.all_gather(allk, torch.tensor([kernel_ms / args.steps], dtype=torch.float64, device="cuda"))
        per_rank_kernel_ms = [float(t.item()) for t in allk]
        dist.all_reduce(red, op=dist.ReduceOp.MAX); dist.all_reduce(tot, op=dist.ReduceOp.SUM)
    total_ms, kernel_ms_max, e2e_s = (float(v) for v in red.cpu())
    useful, executed = (float(v) for v in tot.cpu())

    line = None
    if rank == 0:
        pk = peaks()
        traffic = None
        try:    # DRAM bytes per launch of the dominant kernel, from the committed ncu --set full capture (never measured here)
            traffic = json.load(open(os.path.join(ROOT, "profiles", "traffic.json"))).get(f"{args.workload}:{args.kernel_mode}", {}).get("bytes")
        except Exception:
            pass
        qi = _lib.QueryInfo(); _lib.check(L.fvsrn_query(sn._h, 0, C.byref(qi)))
        value = useful / (total_ms * 1e-3)
        ach = (cnt[0] * w["flop"]) / (kernel_ms * 1e-3) / 1e12            # rank 0's kernel, useful samples only
        frames_per_step = 1 if tiles else world * V
        line = {
            "metric": "srn_samples_per_sec", "value": value, "unit": "samples/s", "n_gpus": world, "steps": args.steps,
            "warmup": max(3, args.warmup), "ms_per_step": total_ms / args.steps, "higher_is_better": True, "scaling": "strong" if tiles else "weak",
            "vs_baseline": None, "dtype": "f16", "data": "synthetic",
            "config": {"workload": f"{args.workload}: {w['desc']}", "views_per_step": frames_per_step, "samples_per_pixel": int(w.get("spp", 1)),
                       "sharding": (f"one frame per step split into {world} contiguous row bands of equal {'estimated' if (args.no_rebalance or world == 1) else 'estimated-then-measured'} march cost {[b[1] for b in bands]}, "
                                    + (f"every rank's kernel writes its band straight into rank 0's frame over NVLink (CUDA IPC peer memory, {fmt}), 4-byte all-reduce as completion signal" if peer_mode else f"bands received in place on rank 0 (NCCL send/recv, {fmt})")) if tiles else
                                   (f"{V} view(s) per GPU, {fmt} pixels written by every rank's kernel straight into rank 0's view array over NVLink (CUDA IPC peer memory), 4-byte all-reduce as completion signal" if peer_mode else
                                    f"{V} view(s) per GPU, {fmt} pixels collected on rank 0 (NCCL {args.collective}), transfer of step k overlaps the render of step k+1" if world > 1 else "single GPU"),
                       "output": f"{fmt} colour written by the DVR kernel (fused extract_color)",
                       "l2": "256 MiB buffer written between timed steps (its device time measured separately and subtracted)",
                       "useful_samples_per_frame": useful / args.steps / frames_per_step, "executed_samples_per_frame": executed / args.steps / frames_per_step,
                       "kernel": {"block_threads": qi.block_threads, "blocks_per_sm": qi.blocks_per_sm, "grid": qi.grid_blocks,
                                  "smem_bytes": qi.smem_bytes_per_block, "regs": qi.regs_per_thread}},
            "ms_per_frame": total_ms / args.steps / (1 if tiles else V),
            "e2e": {"value": useful / e2e_s, "unit": "samples/s", "ms_per_frame": 1e3 * e2e_s / args.steps / (1 if tiles else V),
                    "h2d_bytes_per_step": int(host_cam.numel() * 4 + host_tf.numel() * 4) * (world if world > 1 else 1), "d2h_bytes_per_step": int(d2h_bytes),
                    "what": "fvsrn_render_dvr_host: camera + TF copied up, (B,4,H,W) fp32 colour rendered in four row bands whose D2H copies stream behind the next band's kernel (by_format: rgba8 is written by the kernel straight into the caller's pinned buffer)" if world == 1 else
                            "per step: H2D camera+TF on every rank, render, " + ("peer-memory writes into rank 0's frame" if peer_mode else "NCCL collection on rank 0") + ", D2H of the assembled frames on rank 0",
                    "by_format": e2e_formats or None},
            # one DVR kernel per launch; with per-view keyframe times one launch per view, each preceded by the small kernel that
            # blends the view's keyframe pair (api.cu blend_keyframes; tcgen05 mode)
            "gpu_launches": ((V * (2 if args.kernel_mode == "tcgen05" else 1) if T > 1 else 1) if (not tiles or bh > 0) else 0) * args.steps,
            "roofline": {"bound": "tensor", "achieved": ach, "peak": pk["tflops"], "unit": "TFLOP/s", "frac": ach / pk["tflops"],
                         "traffic": traffic, "kernel": "dvr_tc_kernel (tcgen05.mma + TMEM)" if args.kernel_mode == "tcgen05" else "dvr_kernel (mma.sync)", "flop_per_sample": w["flop"], "peak_source": pk["source"] + ", burst bf16",
                         "executed_frac": (cnt[1] * w["flop"]) / (kernel_ms * 1e-3) / 1e12 / pk["tflops"], "kernel_ms": kernel_ms / args.steps},
            "per_rank_kernel_ms": per_rank_kernel_ms,
            "clocks": clocks, "wall_s_timed_region": t_wall, "batch_query": batch_query,
        }
    # ---- CPU baseline (rank 0, N=1 only): the reference's PyTorch path on the host cores, bounded crop of the same frame
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        import torch_ref
        cores = os.cpu_count() or 1
        torch.set_num_threads(cores)
        tn = torch_ref.TorchNetwork(net)
        kw = dict(width=S, height=S, camera=cam_np[:1], fov_y=scenes.DEFAULT_FOV_Y, stepsize=w["stepsize"], tf=tf_np)   # first view, 1 spp
        torch_ref.render_dvr(tn, **dict(kw, tile=crop_for_cpu(S, 16)))
        t0 = time.perf_counter(); _, ex, _ = torch_ref.render_dvr(tn, **dict(kw, tile=crop_for_cpu(S, 16))); r = ex / (time.perf_counter() - t0)
        px = adaptive_crop(argparse.Namespace(cpu_budget_s=20.0, steps=1, cpu_crop=args.cpu_crop), w, r)
        t0 = time.perf_counter(); _, ex, us = torch_ref.render_dvr(tn, **dict(kw, tile=crop_for_cpu(S, px))); dt = time.perf_counter() - t0
        line["cpu_baseline"] = {"value": us / dt, "unit": "samples/s", "cores": cores, "kind": "port",
                                "sample": f"centre {px}x{px} crop of the {S}^2 frame, oracle/torch_ref.py (reference PyTorch path restated), "
                                          f"{ex} executed / {us} useful samples in {dt:.1f} s"}
        line["speedup_vs_cpu_baseline"] = {"device_timed": line["value"] / (us / dt), "e2e": line["e2e"]["value"] / (us / dt)}
    # ---- the reference's own tensor-core kernel (oracle/_ref PTX, built from /root/reference) on this GPU, same frame: its
    # time (the GPU baseline) and the parity of our frame against it (north_star's gates, measured on the bench frame itself)
    if rank == 0 and world == 1 and not args.no_gpu_reference and not tiles:
        try:
            import ref_kernel
            import parity
            if ref_kernel.available():
                tview = params[0][0].view_times_host and float(keep_alive[0][0]) or 0.0
                rk = ref_kernel.ReferenceDvrKernel(net, time=tview)
                want, ms = rk.render(S, S, cam_np[:1], scenes.DEFAULT_FOV_Y, w["stepsize"], tf_np, timing_iters=3)
                line["reference_gpu_kernel"] = {"ms_per_frame": ms, "value": useful / args.steps / V / (ms * 1e-3), "unit": "samples/s",
                                                "what": "reference ImageEvaluatorSimpleKernel (wmma, NVRTC compute_75 PTX JIT-ed to sm_100), "
                                                        f"block {rk.block}, {rk.blocks_per_sm} block(s)/SM, same weights/camera/TF",
                                                "speedup_of_this_kernel": ms / (kernel_ms / args.steps / V)}
                if int(w.get("spp", 1)) == 1:
                    counted = ref_kernel.ReferenceDvrKernel(net, patched_steps=True, time=tview).render(S, S, cam_np[:1], scenes.DEFAULT_FOV_Y, w["stepsize"], tf_np).cpu().numpy()
                    ie.camera.set_parameters(torch.from_numpy(cam_np[:1])); ie.collect_statistics = True
                    ie.view_times = [tview] if T > 1 else None
                    img = ie.render(S, S).cpu().numpy()
                    par = parity.compare(img, ie.last_step_counts.cpu().numpy(), want.cpu().numpy(), np.rint(counted[:, 4]).astype(np.int32))
                    sn.kernel_mode = "reference_exact"; sn._ensure_prepared()
                    ex_img = ie.render(S, S).cpu().numpy(); ex_steps = ie.last_step_counts.cpu().numpy()
                    sn.kernel_mode = args.kernel_mode
                    wn = want.cpu().numpy()
                    par["reference_exact_mode"] = {"rgba_bit_identical": bool(np.array_equal(ex_img[:, :4], wn[:, :4])),
                                                   "step_counts_bit_identical": bool(np.array_equal(ex_steps, np.rint(counted[:, 4]).astype(np.int32)))}
                    par["gates"] = {"rgba_max_abs": parity.RGBA_TOL, "psnr_db": parity.PSNR_MIN}
                    line["parity"] = par
        except Exception as e:   # no prebuilt PTX for this network shape
            line["reference_gpu_kernel"] = {"unavailable": str(e)[:200]}
    # ---- the other C3 variants the reference's README / benchmark script name (SnakeAlt:2 activations, Gaussian Fourier
    # matrix): kernel time of the same frame, device-timed, 5 steps each
    if rank == 0 and world == 1 and args.workload == "C3" and not args.no_variants:
        line["variants"] = {}
        for name, kwn in (("C3-snake (SnakeAlt:2, NeRF Fourier)", dict(WORKLOADS["C3-snake"]["net"])),
                          ("C3-gauss (ReLU, Gaussian Fourier std 1)", dict(WORKLOADS["C3"]["net"], fourier="gauss", fourier_std=1.0)),
                          ("C3-snake-gauss (SnakeAlt:2, Gaussian Fourier std 1)", dict(WORKLOADS["C3-snake"]["net"], fourier="gauss", fourier_std=1.0))):
            vn = scenes.make_network(**kwn)
            vs = pr.SceneNetwork.from_bytes(volnet.dumps(vn)); vs.kernel_mode = args.kernel_mode
            vv = pr.VolumeInterpolationNetwork(); vv.set_network(vs); vs._ensure_prepared()
            cnt2 = torch.zeros(2, dtype=torch.int64, device="cuda")
            pv = params[0][0]; pv.counters_dev = cnt2.data_ptr()
            for _ in range(2): _lib.check(L.fvsrn_render_dvr(vs._h, C.byref(pv), params[0][1], stream))
            cnt2.zero_()
            ve = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
            ms = 0.0
            for _ in range(5):
                flush.zero_(); ve[0].record(); _lib.check(L.fvsrn_render_dvr(vs._h, C.byref(pv), params[0][1], stream)); ve[1].record()
                torch.cuda.synchronize(); ms += ve[0].elapsed_time(ve[1])
            us = float(cnt2.cpu()[0]) / 5
            line["variants"][name] = {"ms_per_frame": ms / 5, "useful_samples_per_frame": us, "value": us / (ms / 5 * 1e-3),
                                      "tensor_frac_of_peak": us * w["flop"] / (ms / 5 * 1e-3) / 1e12 / peaks()["tflops"]}
            pv.counters_dev = counters.data_ptr()
    if rank == 0:
        _emit(line)
    if peer is not None:
        frames = gathered = params = res0 = None
        peer.close()
    if world > 1:
        dist.barrier(); dist.destroy_process_group()
    return 0


_REAL_STDOUT = None


def _guard_stdout():
    """The contract is ONE JSON line on stdout. Libraries print there too (NCCL's version banner at communicator
    creation): everything except the result line is sent to stderr."""
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)


def _emit(line):
    sys.stdout.flush()
    os.write(_REAL_STDOUT if _REAL_STDOUT is not None else 1, (json.dumps(line) + "\n").encode())


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="C3", choices=sorted(WORKLOADS))
    ap.add_argument("--cpu-crop", type=int, default=256, help="max edge of the centre crop the CPU legs render")
    ap.add_argument("--cpu-budget-s", type=float, default=120.0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-batch-query", action="store_true", help="skip the C2 world-space batch query measurement")
    ap.add_argument("--no-gpu-reference", action="store_true", help="skip timing the reference's own kernel (oracle/_ref PTX) and the parity check against it")
    ap.add_argument("--no-variants", action="store_true", help="skip the C3 SnakeAlt / Gaussian-Fourier variants")
    ap.add_argument("--no-rebalance", action="store_true", help="--sharding tiles: keep the box-only cost estimate (no measured feedback)")
    ap.add_argument("--sharding", default="views", choices=["views", "tiles"],
                    help="N GPUs: views = every rank renders its own views per step (weak scaling); tiles = ONE frame per step split "
                         "into cost-balanced row bands (strong scaling, BASELINE C4)")
    ap.add_argument("--gather-format", default="rgba8", choices=["rgba8", "rgba_f16", "rgba_f32"],
                    help="pixel format the DVR kernel writes and NCCL collects (fused colour extraction)")
    ap.add_argument("--collective", default="gather", choices=["gather", "all_gather", "peer"],
                    help="N>1: how the finished pixels reach rank 0. gather / all_gather: NCCL collective behind the render (fv-srn_b200/dist.py "
                         "gather_views / gather_row_bands); peer: the DVR kernel of every rank writes its pixels straight into rank 0's frame "
                         "over NVLink (dist.PeerFrames, CUDA IPC), a 4-byte all-reduce is the completion signal")
    ap.add_argument("--kernel-mode", default="tcgen05", choices=["tcgen05", "mma_sync"],
                    help="tcgen05: tcgen05.mma/TMEM kernel (default); mma_sync: reference-precision fp16-accumulate kernel")
    args = ap.parse_args()
    _guard_stdout()
    if args.impl == "reference":
        return run_reference(args)
    return run_b200(args)


if __name__ == "__main__":
    sys.exit(main())
