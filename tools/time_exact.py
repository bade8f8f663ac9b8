"""Builder diagnostic: ms per C3 frame of each kernel mode (CUDA events, L2 flushed), incl. the reference-exact validation mode."""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import fvsrn_b200  # noqa
from fvsrn_b200 import pyrenderer as pr, scenes, volnet
plume = np.load(os.path.join(ROOT, "tests", "golden", "plume_tf.npy"))
net = scenes.make_network(hidden=32, num_layers=4, grid_res=32, seed=42, activation="ReLU", last_gain=6.0)
cam = scenes.camera_reference_frame(**scenes.DEFAULT_CAMERA)
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
for mode in ("tcgen05", "mma_sync", "reference_exact"):
    sn = pr.SceneNetwork.from_bytes(volnet.dumps(net)); sn.kernel_mode = mode
    vol = pr.VolumeInterpolationNetwork(); vol.set_network(sn)
    ie = pr.ImageEvaluatorSimple(); ie.volume = vol
    ie.camera.set_parameters(torch.from_numpy(np.ascontiguousarray(cam, np.float32)))
    ie.ray_evaluator.tf.tensor.value = torch.from_numpy(plume).cuda(); ie.ray_evaluator.stepsize = 1 / 512
    out = ie.render(1024, 1024); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ms = []
    for _ in range(5):
        flush.zero_(); e0.record(); ie.render(1024, 1024, out); e1.record(); torch.cuda.synchronize(); ms.append(e0.elapsed_time(e1))
    print(f"C3 1024^2 step 1/512, kernel mode {mode}: {np.median(ms):.3f} ms/frame")
