"""
Two or more GPUs, one process per GPU (torch.distributed.run): the peer-memory frame path (dist.PeerFrames - every rank's DVR
kernel writes its pixels straight into rank 0's frame over NVLink) against the NCCL gather path (dist.gather_views /
gather_row_bands) and against rank 0 rendering everything alone. Bit-identical pixels or exit code 1.
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 tools/check_peer_frames.py
"""
import ctypes as C
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch
    import torch.distributed as dist
    import fvsrn_b200  # noqa: F401
    from fvsrn_b200 import pyrenderer as pr, scenes, volnet, _lib, dist as fdist
    world = int(os.environ["WORLD_SIZE"]); rank = int(os.environ["RANK"]); local = int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    L = _lib.lib()
    S, V, step = 320, 2, 1 / 128
    net = scenes.make_network(hidden=32, num_layers=4, grid_res=32, seed=42, activation="ReLU", last_gain=6.0)
    sn = pr.SceneNetwork.from_bytes(volnet.dumps(net))
    vol = pr.VolumeInterpolationNetwork(); vol.set_network(sn)
    ie = pr.ImageEvaluatorSimple(); ie.volume = vol
    ie.ray_evaluator.tf.tensor.value = torch.from_numpy(np.load(os.path.join(ROOT, "tests", "golden", "plume_tf.npy"))).cuda()
    ie.ray_evaluator.stepsize = step
    sn._ensure_prepared()
    tf_dev = ie.ray_evaluator.tf._tensor()
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    cams = scenes.camera_sweep(world * V)
    failures = []

    def render(cam_np, base, strides, rect=None):
        ie.camera.set_parameters(torch.from_numpy(np.ascontiguousarray(cam_np)))
        cam_dev = ie.camera.get_parameters()
        p = ie._params(S, S, cam_dev, tf_dev, strides)
        p.output_format, p.channel_mode = pr.OUTPUT_FORMATS["rgba8"], 3
        if rect: p.tile_x0, p.tile_y0, p.tile_w, p.tile_h = rect
        _lib.check(L.fvsrn_render_dvr(sn._h, C.byref(p), base, stream))
        return cam_dev      # keep the camera tensor alive until the launch has been enqueued behind it

    strides = (S * S, 0, S, 1)
    # ---- view sharding: views [rank V, (rank + 1) V)
    alone = torch.zeros((world * V, S, S), dtype=torch.int32, device="cuda")
    if rank == 0:
        render(cams, alone.data_ptr(), strides)
    mine = torch.zeros((V, S, S), dtype=torch.int32, device="cuda")
    render(cams[rank * V:(rank + 1) * V], mine.data_ptr(), strides)
    gathered = fdist.gather_views(mine, world * V)
    peer = fdist.PeerFrames((world * V, S, S), torch.int32, dst=0, slots=2)
    for k in range(4):          # both slots, twice
        slot = k & 1
        render(cams[rank * V:(rank + 1) * V], peer.address(slot, rank * V), strides)
        wk = peer.signal(slot); wk.wait()
        torch.cuda.synchronize()
        if rank == 0:
            if not torch.equal(peer.tensor(slot), alone): failures.append(f"views: peer frame of step {k} differs from rank 0 rendering alone")
            if not torch.equal(peer.tensor(slot), gathered): failures.append(f"views: peer frame of step {k} differs from the NCCL gather")
            peer.tensor(slot).zero_()
        dist.barrier()
    if rank == 0 and not alone.any(): failures.append("views: empty frames")
    peer.close()

    # ---- band sharding of one frame: rows [y0, y0 + bh)
    cam1 = cams[0:1]
    bands = fdist.partition_rows_by_cost(scenes.estimate_row_cost(cam1, scenes.DEFAULT_FOV_Y, S, S, net.box_min, net.box_size), world, align=8)
    y0, bh = bands[rank]
    full = torch.zeros((1, S, S), dtype=torch.int32, device="cuda")
    if rank == 0:
        render(cam1, full.data_ptr(), strides)
    peer = fdist.PeerFrames((1, S, S), torch.int32, dst=0, slots=1)
    if bh > 0:
        render(cam1, peer.address(0), strides, rect=(0, y0, S, bh))
    peer.signal(0).wait(); torch.cuda.synchronize()
    if rank == 0 and not torch.equal(peer.tensor(0), full): failures.append(f"bands {bands}: peer frame differs from rank 0 rendering alone")
    peer.close()

    bad = torch.tensor([len(failures)], device="cuda"); dist.all_reduce(bad)
    if rank == 0:
        print("peer frames:", "OK" if not failures else failures, f"(world {world}, {world * V} views of {S}^2, bands {bands})", flush=True)
    dist.barrier(); dist.destroy_process_group()
    sys.exit(1 if int(bad.item()) else 0)


if __name__ == "__main__":
    main()
