"""Per-stage device times of forward+backward(+Adam) on a BASELINE config (CUDA events inside the library)."""
import argparse, importlib, json, os, sys, time
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
pkg = importlib.import_module("3d_gaussian_splatting_b200")
syn = importlib.import_module("3d_gaussian_splatting_b200.synthetic")

ap = argparse.ArgumentParser()
ap.add_argument("--config", default="cfg2_1p6m_1246x825_sh3")
ap.add_argument("--iters", type=int, default=10)
ap.add_argument("--N", type=int, default=None)
a = ap.parse_args()
t0 = time.time()
scene, cams, deg = syn.make_config(a.config, n_cameras=8, N=a.N)
sc, rot, op, shs = scene.activated()
dev = lambda x: torch.from_numpy(np.ascontiguousarray(x)).cuda()
xyz, sc, rot, op, shs = dev(scene.xyz), dev(sc), dev(rot), dev(op), dev(shs)
print("scene built", time.time() - t0, flush=True)
rast = pkg.Rasterizer(0)
rast.set_profiling(True)
H, W = cams[0].image_height, cams[0].image_width
dL = (torch.rand((3, H, W), device="cuda") * 2 - 1) / (H * W)
acc = {}
for it in range(a.iters + 3):
    cam = cams[it % len(cams)]
    sg = pkg.GaussianRasterizationSettings.from_camera(cam, (0, 0, 0), deg, deg)
    color, radii = rast.forward(sg, xyz, sc, rot, op, shs=shs)
    g = rast.backward(dL)
    torch.cuda.synchronize()
    ms = rast.stage_ms()
    if it >= 3:
        for k, v in ms.items():
            acc.setdefault(k, []).append(v)
D = int(rast.debug_get("NUM_RENDERED")[0])
nvis = int((radii > 0).sum().item())
ncontrib = rast.debug_get("N_CONTRIB")
res = {k: float(np.median(v)) for k, v in acc.items()}
res["total_ms"] = sum(res.values())
print(json.dumps(dict(config=a.config, N=scene.N, N_vis=nvis, D=D, mean_n_contrib=float(ncontrib.mean()),
                      max_n_contrib=int(ncontrib.max()), stage_ms=res)))
