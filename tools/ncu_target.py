"""Builder tool: a short program for `ncu --set full -k regex:...`: two C3 frames (dvr_tc_kernel, SPEC 1) and two C2 batch
queries (eval_tc_kernel) through the product API."""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import fvsrn_b200  # noqa
from fvsrn_b200 import pyrenderer as pr, scenes, volnet
plume = np.load(os.path.join(ROOT, "tests", "golden", "plume_tf.npy"))
net = scenes.make_network(hidden=32, num_layers=4, grid_res=32, seed=42, activation="ReLU", last_gain=6.0)
sn = pr.SceneNetwork.from_bytes(volnet.dumps(net))
vol = pr.VolumeInterpolationNetwork(); vol.set_network(sn)
ie = pr.ImageEvaluatorSimple(); ie.volume = vol
ie.camera.set_parameters(torch.from_numpy(np.ascontiguousarray(scenes.camera_reference_frame(**scenes.DEFAULT_CAMERA), np.float32)))
ie.ray_evaluator.tf.tensor.value = torch.from_numpy(plume).cuda(); ie.ray_evaluator.stepsize = 1 / 512
g = torch.Generator(device="cuda"); g.manual_seed(1)
pos = torch.rand((1 << 24, 3), device="cuda", generator=g)
for _ in range(2):
    img = ie.render(1024, 1024)
    out = vol.evaluate(pos)
torch.cuda.synchronize()
print("ok", float(img[:, 3].mean()), float(out.mean()))
