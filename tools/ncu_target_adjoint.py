"""Builder tool: a short program for `ncu --set full -k regex:dvr_exact`: one 512^2 C3 frame with Phong shading in adjoint gradient
mode (dvr_exact_kernel<.., ADJ = true>) through the product API; prints the fraction of warp-steps that shade something."""
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
vol.gradient_mode = pr.VolumeInterpolationNetwork.ADJOINT_METHOD
ie = pr.ImageEvaluatorSimple(); ie.volume = vol
ie.camera.set_parameters(torch.from_numpy(np.ascontiguousarray(scenes.camera_reference_frame(**scenes.DEFAULT_CAMERA), np.float32)))
ie.ray_evaluator.tf.tensor.value = torch.from_numpy(plume).cuda(); ie.ray_evaluator.stepsize = 1 / 512
b = ie.ray_evaluator.brdf; b.enable_phong = True; b.light_follows_camera = False; b.light_direction = pr.double3(0.3, -0.5, -0.8)
img = ie.render(512, 512)
torch.cuda.synchronize()
pos = torch.rand((1 << 20, 3), device="cuda")
d = vol.evaluate(pos)
print("ok", float(img[:, 3].mean()), "fraction of random positions with density >= 0:", float((d >= 0).float().mean()))
