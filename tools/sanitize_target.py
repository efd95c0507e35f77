"""Small end-to-end exercise of every single-GPU kernel for compute-sanitizer (tools/gpu_sanitize.sh): forward and
backward on the raw-parameter path (TMA-staged K1 / K4b, one-sweep sorts, both blend kernels, D-SSIM loss), the generic
kernels (get_features layout, ragged N), the fused backward + Adam step, the separate Adam + statistics kernel,
densification (plan / apply / opacity reset) and a second step at the new size.  BASELINE.json configs[0]-like sizes."""
import importlib
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
gm = importlib.import_module("3d_gaussian_splatting_b200.gaussian_model")
syn = importlib.import_module("3d_gaussian_splatting_b200.synthetic")
cammod = importlib.import_module("3d_gaussian_splatting_b200.camera")

scene = syn.make_scene(3000 + 37, 2, seed=7)
scene.log_scales += np.float32(0.5)
cams = [cammod.look_at_camera((0.5 * np.cos(a), 0.3, 3.5), (0, 0, 0), (0, 1, 0), 120.0, 75, 120) for a in (0.0, 0.7)]
model = gm.GaussianModel(2, 0).load_scene(scene)
opt = gm.OptimizationParams()
opt.densify_from_iter, opt.densification_interval, opt.densify_grad_threshold = 1, 2, 1e-7
tr = gm.Trainer(model, opt, lambda_dssim=0.2)
gt_model = gm.GaussianModel(2, 0).load_scene(scene)
gt_model.features_dc_ += 0.1
gts = [gm.render(c, gt_model, gm.PipelineParams(), (0., 0., 0.), tr.rast)["render"].cpu().pin_memory() for c in cams]
tr.reserve_pairs(400_000)
losses = tr.train(cams, gts, 4, densify_extent=3.0)           # fused steps, a densification, steps at the new N
tr.fused_step = False
losses.append(tr.train_step(cams[0], gts[0]))                 # separate K4b + K5
# generic kernels: get_features layout, activated inputs, precomputed colours
g = tr.g
st = gm.GaussianRasterizationSettings.from_camera(cams[1], (0.1, 0.2, 0.3), 2, 2)
img, radii = tr.rast.forward(st, g.get_xyz(), g.get_scaling(), g.get_rotation(), g.get_opacity(), shs=g.get_features())
tr.rast.backward(torch.rand_like(img) / img.numel())
torch.cuda.synchronize()
print("sanitize target ok: N", g.N, "losses", [round(float(l), 5) for l in losses])
