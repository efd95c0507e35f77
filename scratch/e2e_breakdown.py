import importlib, sys, time
import numpy as np, torch
sys.path.insert(0, "/root/repo")
syn = importlib.import_module("3d_gaussian_splatting_b200.synthetic")
gm = importlib.import_module("3d_gaussian_splatting_b200.gaussian_model")
scene, cams, deg = syn.make_config("cfg2_1p6m_1246x825_sh3", n_cameras=8)
model = gm.GaussianModel(deg, 0).load_scene(scene)
tr = gm.Trainer(model, gm.OptimizationParams(), lambda_dssim=0.2)
gts = []
for c in cams:
    out = gm.render(c, model, gm.PipelineParams(), (0., 0., 0.), tr.rast)
    gts.append(out["render"].cpu().pin_memory())
torch.cuda.synchronize()
tr.reserve_pairs(8_000_000)
T = {}
def tick(name, t0):
    t = time.perf_counter(); T.setdefault(name, []).append((t - t0) * 1e6); return t
rast = tr.rast
orig_fwd, orig_loss, orig_bwd = rast.forward, rast.loss, rast.backward_adam_step
def wrap(fn, name):
    def w(*a, **k):
        t0 = time.perf_counter(); r = fn(*a, **k); tick(name, t0); return r
    return w
rast.forward = wrap(orig_fwd, "forward_call"); rast.loss = wrap(orig_loss, "loss_call"); rast.backward_adam_step = wrap(orig_bwd, "bwd_adam_call")
for i in range(60):
    t0 = time.perf_counter()
    l = tr.train_step(cams[i % 8], gts[i % 8])
    tick("train_step_total", t0)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
torch.cuda.synchronize(); t0 = time.perf_counter(); e0.record()
for i in range(40):
    tr.train_step(cams[i % 8], gts[i % 8])
e1.record(); torch.cuda.synchronize()
print("e2e ms/step wall", (time.perf_counter() - t0) / 40 * 1e3, "events", e0.elapsed_time(e1) / 40)
# device-only for the same loop
gd = [g.cuda() for g in gts]
torch.cuda.synchronize(); e0.record()
for i in range(40):
    tr.train_step_device(cams[i % 8], gd[i % 8])
e1.record(); torch.cuda.synchronize()
print("device ms/step", e0.elapsed_time(e1) / 40)
for k, v in T.items():
    print(f"{k:20s} median {np.median(v[10:]):8.1f} us")
