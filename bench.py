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
  N GPUs   weak scaling: every rank renders its own view of a camera sweep per step (view-sharded multi-view batch);
           rank 0 gathers the RGBA tiles over NCCL inside the timed step.
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
    sn = pr.SceneNetwork.from_bytes(volnet.dumps(net))
    sn.kernel_mode = args.kernel_mode
    vol = pr.VolumeInterpolationNetwork(); vol.set_network(sn)
    ie = pr.ImageEvaluatorSimple(); ie.volume = vol
    V = int(w.get("views", 1))                              # views per rank per step (view-sharded multi-view batch)
    cams = scenes.camera_sweep(world * V)
    cam_np = np.ascontiguousarray(cams[rank * V:(rank + 1) * V])
    ie.camera.set_parameters(torch.from_numpy(cam_np))
    ie.samples_per_pixel = int(w.get("spp", 1))
    keep_alive = []
    T = int(w["net"].get("time_keyframes", 0))
    ie.ray_evaluator.tf.tensor.value = torch.from_numpy(tf_np).cuda()
    ie.ray_evaluator.stepsize = w["stepsize"]
    sn._ensure_prepared()
    out = torch.empty((V, 8, S, S), dtype=torch.float32, device="cuda")
    counters = torch.zeros(2, dtype=torch.int64, device="cuda")
    cam_dev = ie.camera.get_parameters(); tf_dev = ie.ray_evaluator.tf._tensor()
    p = ie._params(S, S, cam_dev, tf_dev, out.stride())
    p.counters_dev = counters.data_ptr()
    if T > 1:   # per-view time: a golden-ratio sequence over the keyframe range, so that every rank's views need the temporal
        # interpolation of two keyframe grids (the expensive case) in the same proportion at every N
        times = [(T - 1) * (((rank * V + i) * 0.6180339887) % 1.0) for i in range(V)]
        vt = (C.c_float * V)(*times); keep_alive.append(vt)
        p.view_times_host = C.cast(vt, C.c_void_p)
    gathered = torch.empty((world * V, 4, S, S), dtype=torch.float32, device="cuda") if world > 1 else None
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")   # > 126 MB L2

    def step(ev=None):
        if ev: ev[0].record()
        _lib.check(L.fvsrn_render_dvr(sn._h, C.byref(p), out.data_ptr(), stream))
        if ev: ev[1].record()
        rgba = ie.extract_color(out)
        if world > 1:
            fdist.gather_views(rgba, world * V, out=gathered, collective=args.collective)  # one NCCL collective: the finished RGBA views land on rank 0
        if ev: ev[2].record()
        return rgba

    def barrier():
        if world > 1: dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(3, args.warmup)):
        step(); flush.zero_()
    barrier()
    counters.zero_()
    evs = [[torch.cuda.Event(enable_timing=True) for _ in range(3)] for _ in range(args.steps)]
    sampler = ClockSampler(local) if rank == 0 else None
    if sampler: sampler.start()
    barrier()
    t_wall = time.perf_counter()
    for k in range(args.steps):
        step(evs[k]); flush.zero_()
    barrier()
    t_wall = time.perf_counter() - t_wall
    total_ms = sum(e[0].elapsed_time(e[2]) for e in evs)
    kernel_ms = sum(e[0].elapsed_time(e[1]) for e in evs)
    cnt = counters.cpu().numpy().astype(np.float64)          # [useful, executed] over the K timed steps

    # ---- e2e: the reference-facing call with HOST buffers (camera + TF in, (1,8,S,S) image out)
    host_img = torch.empty((V, 8, S, S), dtype=torch.float32).pin_memory()
    host_cam = torch.from_numpy(cam_np.copy()).pin_memory(); host_tf = torch.from_numpy(np.ascontiguousarray(tf_np)).pin_memory()
    q = ie._params(S, S, host_cam, host_tf, (8 * S * S, S * S, S, 1))
    q.camera_matrix_dev = host_cam.data_ptr(); q.tf_dev = host_tf.data_ptr()
    q.view_times_host = p.view_times_host

    def e2e_step():
        _lib.check(L.fvsrn_render_dvr_host(sn._h, C.byref(q), host_img.data_ptr(), stream))
        return float(host_img[0, 3, S // 2, S // 2])        # read the result on the host

    for _ in range(2): e2e_step()
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps): e2e_step(); flush.zero_()
    torch.cuda.synchronize()
    e2e_s = time.perf_counter() - t0
    flush_t0 = time.perf_counter()
    for _ in range(args.steps): flush.zero_()
    torch.cuda.synchronize()
    e2e_s = max(1e-9, e2e_s - (time.perf_counter() - flush_t0))   # the L2 flush is hygiene, not part of the call
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
        del pos, dens

    red = torch.tensor([total_ms, kernel_ms, e2e_s], dtype=torch.float64, device="cuda")
    tot = torch.tensor([cnt[0], cnt[1]], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(red, op=dist.ReduceOp.MAX); dist.all_reduce(tot, op=dist.ReduceOp.SUM)
    total_ms, kernel_ms, e2e_s = (float(v) for v in red.cpu())
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
        line = {
            "metric": "srn_samples_per_sec", "value": value, "unit": "samples/s", "n_gpus": world, "steps": args.steps,
            "warmup": max(3, args.warmup), "ms_per_step": total_ms / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f16", "data": "synthetic",
            "config": {"workload": f"{args.workload}: {w['desc']}", "views_per_step": world * V, "samples_per_pixel": int(w.get("spp", 1)), "sharding": f"{V} view(s) per GPU, RGBA collected on rank 0 (NCCL {args.collective})" if world > 1 else "single GPU",
                       "l2": "256 MiB buffer written between timed steps (outside the per-step CUDA-event pairs)",
                       "useful_samples_per_frame": useful / args.steps / world / V, "executed_samples_per_frame": executed / args.steps / world / V,
                       "kernel": {"block_threads": qi.block_threads, "blocks_per_sm": qi.blocks_per_sm, "grid": qi.grid_blocks,
                                  "smem_bytes": qi.smem_bytes_per_block, "regs": qi.regs_per_thread}},
            "ms_per_frame": total_ms / args.steps / V,
            "e2e": {"value": useful / e2e_s, "unit": "samples/s", "ms_per_frame": 1e3 * e2e_s / args.steps / V,
                    "h2d_bytes_per_step": int(host_cam.numel() * 4 + host_tf.numel() * 4), "d2h_bytes_per_step": int(host_img.numel() * 4)},
            "gpu_launches": (V if T > 1 else 1) * args.steps + args.steps,
            "roofline": {"bound": "tensor", "achieved": ach, "peak": pk["tflops"], "unit": "TFLOP/s", "frac": ach / pk["tflops"],
                         "traffic": traffic, "kernel": "dvr_tc_kernel (tcgen05.mma + TMEM)" if args.kernel_mode == "tcgen05" else "dvr_kernel (mma.sync)", "flop_per_sample": w["flop"], "peak_source": pk["source"] + ", burst bf16",
                         "executed_frac": (cnt[1] * w["flop"]) / (kernel_ms * 1e-3) / 1e12 / pk["tflops"], "kernel_ms": kernel_ms / args.steps},
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
    # ---- the reference's own tensor-core kernel (oracle/_ref PTX, built from /root/reference) on this GPU, same frame
    if rank == 0 and world == 1 and not args.no_gpu_reference:
        try:
            import ref_kernel
            if ref_kernel.available():
                rk = ref_kernel.ReferenceDvrKernel(net)
                _, ms = rk.render(S, S, cam_np[:1], scenes.DEFAULT_FOV_Y, w["stepsize"], tf_np, timing_iters=3)
                line["reference_gpu_kernel"] = {"ms_per_frame": ms, "value": useful / args.steps / V / (ms * 1e-3), "unit": "samples/s",
                                                "what": "reference ImageEvaluatorSimpleKernel (wmma, NVRTC compute_75 PTX JIT-ed to sm_100), "
                                                        f"block {rk.block}, {rk.blocks_per_sm} block(s)/SM, same weights/camera/TF",
                                                "speedup_of_this_kernel": ms / (kernel_ms / args.steps / V)}
        except Exception as e:   # no prebuilt PTX for this network shape
            line["reference_gpu_kernel"] = {"unavailable": str(e)[:200]}
    if rank == 0:
        _emit(line)
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
    ap.add_argument("--no-gpu-reference", action="store_true", help="skip timing the reference's own kernel (oracle/_ref PTX)")
    ap.add_argument("--collective", default="gather", choices=["gather", "all_gather"],
                    help="N>1: how the finished RGBA views reach rank 0 (fv-srn_b200/dist.py gather_views)")
    ap.add_argument("--kernel-mode", default="tcgen05", choices=["tcgen05", "mma_sync"],
                    help="tcgen05: tcgen05.mma/TMEM kernel (default); mma_sync: reference-precision fp16-accumulate kernel")
    args = ap.parse_args()
    _guard_stdout()
    if args.impl == "reference":
        return run_reference(args)
    return run_b200(args)


if __name__ == "__main__":
    sys.exit(main())
