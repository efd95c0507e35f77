import importlib, sys, os
import numpy as np, torch
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/tests")
from helpers import gpu_settings, small_scene
pkg = importlib.import_module("3d_gaussian_splatting_b200")
L = importlib.import_module("3d_gaussian_splatting_b200._lib")
rast = pkg.Rasterizer(0)
dev = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
N, deg = 3000, 2
scene, cam = small_scene(N=N, H=96, W=128, deg=deg, seed=43, scale_boost=0.5)
K = (deg + 1) ** 2
widths = [3, 3, 3 * (K - 1), 1, 3, 4]
sg = gpu_settings(cam, deg, deg)
raw = [scene.xyz, scene.features_dc, scene.features_rest, scene.opacity_logits, scene.log_scales, scene.raw_rotations]
P = [dev(np.ascontiguousarray(a, dtype=np.float32)) for a in raw]
lr = [1.6e-4, 2.5e-3, 1.25e-4, 0.05, 5e-3, 1e-3]
rng = np.random.default_rng(2)
M = [dev((rng.standard_normal(p.shape) * 1e-4).astype(np.float32)) for p in P]
V = [dev((rng.random(p.shape) * 1e-8).astype(np.float32)) for p in P]
maxr = torch.zeros(N, dtype=torch.int32, device="cuda"); acc = torch.zeros(N, device="cuda"); den = torch.zeros(N, device="cuda")
rs = np.random.default_rng(5)
dL = dev(((rs.random((3, 96, 128)) * 2 - 1) / (96 * 128)).astype(np.float32))
_, radii = rast.forward(sg, P[0], P[4], P[5], P[3], sh_dc=P[1], sh_rest=P[2], raw_params=True)
G = rast.backward(dL)
A, MA, VA = [p.clone() for p in P], [m.clone() for m in M], [v.clone() for v in V]
Gk = [G["means3D"], G["sh_dc"], G.get("sh_rest"), G["opacities"], G["scales"], G["rotations"]]
rast.adam_densify_step(N, A, Gk, MA, VA, widths, lr, [1] * 6)
L.check(rast._ctx, L.lib.gs_debug_reuse_blend_gradients(rast._ctx, 1))
rast.backward_adam_step(dL, M, V, lr, [1] * 6, max_radii2D=maxr, xyz_grad_accum=acc, denom=den)
torch.cuda.synchronize()
for k in range(6):
    for name, x, y in (("p", P[k], A[k]), ("m", M[k], MA[k]), ("v", V[k], VA[k])):
        d = (x != y).reshape(N, -1)
        rows = d.any(dim=1).nonzero().flatten()
        print(k, name, int(d.sum()), "rows", rows[:6].tolist(), rows[-3:].tolist() if len(rows) else "",
              float((x - y).abs().max()))
