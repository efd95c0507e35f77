"""Checkpoint / export formats next to the parameter store.

capture() / restore() mirror GaussianModel::CoreParams::capture / restore
(src/scene/gaussian_model.cpp:76-202): ONE tensors file holding the same eleven entries in the same order
(active_sh_degree, xyz, features_dc, features_rest, scaling, rotation, opacity, max_radii2D,
xyz_gradient_accum, denom, spatial_lr_scale; :85-97) plus SIX optimiser files (xyz, f_dc, f_rest, opacity,
scaling, rotation; :100-131), each restored only if it exists (:163-201).

Two containers, selected by `fmt`:

"libtorch" (default) -- the reference's own: the archives torch::save writes and torch::load reads.  The tensors
file is what `torch::save(std::vector<torch::Tensor>, path)` produces (a TorchScript module whose attributes "0" ..
"10" are the tensors); an optimiser file is what `torch::save(torch::optim::Adam&, path)` produces (attributes
pytorch_version, state/<key>/{step, exp_avg, exp_avg_sq}, param_groups/...options{lr, betas, eps, weight_decay,
amsgrad}).  They are built here through torch's script-module API and read back with torch.jit.load, so both sides
can exchange checkpoints: tests/test_checkpoint.py loads the files written here with libtorch's torch::load
(tests/ckpt/ckpt_module.cpp) and reads files written by libtorch's torch::save.

"flat" -- a dependency-free little-endian dump, the container the NATIVE trainer (csrc/trainer.cu, gs_trainer_save /
gs_trainer_load) reads and writes, since libgsplat_b200.so does not link libtorch:

    file   := magic "GSB200CK" | u32 version (1) | u32 n_entries | entry*
    entry  := u8 dtype (0 f32, 1 i32, 2 i64, 3 f64) | u8 ndim | u16 0 | u64 dims[ndim] | raw C-order bytes

An optimiser file holds four entries: step (i64 scalar), lr (f64 scalar), exp_avg, exp_avg_sq -- the state
torch::optim::Adam keeps per parameter.  search_for_max_iteration mirrors src/utils/system_utils.cpp:7-31.
save_ply writes the point-cloud layout of the published 3DGS trainer (x y z nx ny nz f_dc_* f_rest_* opacity
scale_* rot_*, binary little-endian, f_rest channel-major) that viewers of the reference's outputs expect.
"""
from __future__ import annotations

import os
import struct
from typing import Sequence

import numpy as np
import torch

MAGIC = b"GSB200CK"
VERSION = 1
_DT = {np.dtype(np.float32): 0, np.dtype(np.int32): 1, np.dtype(np.int64): 2, np.dtype(np.float64): 3}
_DT_INV = {v: k for k, v in _DT.items()}
GROUPS = ("xyz", "f_dc", "f_rest", "opacity", "scaling", "rotation")   # GaussianModel::setup order


def write_entries(path: str, arrays: Sequence[np.ndarray]) -> None:
    tmp = path + ".tmp"
    with open(tmp, "wb") as f:
        f.write(MAGIC + struct.pack("<II", VERSION, len(arrays)))
        for a in arrays:
            a = np.asarray(a, order="C")    # (ascontiguousarray would promote 0-d scalars to 1-d)
            if a.dtype not in _DT:
                raise TypeError(f"unsupported dtype {a.dtype}")
            f.write(struct.pack("<BBH", _DT[a.dtype], a.ndim, 0))
            f.write(struct.pack(f"<{a.ndim}Q", *a.shape))
            f.write(a.tobytes())
    os.replace(tmp, path)    # a crash never leaves a truncated checkpoint under the final name


def read_entries(path: str):
    with open(path, "rb") as f:
        head = f.read(16)
        if len(head) < 16 or head[:8] != MAGIC:
            raise ValueError(f"{path}: not a gsplat_b200 checkpoint")
        version, n = struct.unpack("<II", head[8:])
        if version != VERSION:
            raise ValueError(f"{path}: checkpoint version {version}, this build reads {VERSION}")
        out = []
        for _ in range(n):
            dt, ndim, _pad = struct.unpack("<BBH", f.read(4))
            dims = struct.unpack(f"<{ndim}Q", f.read(8 * ndim))
            dtype = _DT_INV[dt]
            count = int(np.prod(dims)) if ndim else 1
            buf = f.read(count * dtype.itemsize)
            if len(buf) != count * dtype.itemsize:
                raise ValueError(f"{path}: truncated")
            out.append(np.frombuffer(buf, dtype=dtype).reshape(dims).copy())
        return out


def _np(t: torch.Tensor) -> np.ndarray:
    return t.detach().cpu().numpy()


# ---- libtorch archives (the reference's container) ---------------------------------------------------------------
def _new_module_factory():
    cu = torch._C.CompilationUnit()
    return lambda: torch._C.ScriptModule("__torch__.Module", cu, True)


def _save_module(mod, path: str) -> None:
    tmp = path + ".tmp"
    mod.save(tmp)
    os.replace(tmp, path)


def _jit_load(path: str):
    import warnings
    try:
        with warnings.catch_warnings():
            warnings.simplefilter("ignore")
            return torch.jit.load(path, map_location="cpu")
    except RuntimeError as e:      # truncated / overwritten file: same exception type as the flat reader
        raise ValueError(f"{path}: not a readable checkpoint ({str(e).splitlines()[0]})") from None


def write_tensor_archive(path: str, tensors: Sequence[torch.Tensor]) -> None:
    """The file `torch::save(std::vector<torch::Tensor>{...}, path)` writes: tensor i under the key str(i)."""
    new = _new_module_factory()
    m = new()
    for i, t in enumerate(tensors):
        m._register_attribute(str(i), torch._C.TensorType.get(), t.detach().cpu().contiguous())
    _save_module(m, path)


def read_tensor_archive(path: str):
    """Inverse (also reads archives written by libtorch): the tensors under the keys "0", "1", ... in order."""
    m = _jit_load(path)
    out = []
    while hasattr(m, str(len(out))):
        out.append(getattr(m, str(len(out))).detach())
    if not out:
        raise ValueError(f"{path}: not a torch::save(std::vector<torch::Tensor>) archive")
    return out


def write_adam_archive(path: str, step: int, lr: float, exp_avg: torch.Tensor, exp_avg_sq: torch.Tensor,
                       betas=(0.9, 0.999), eps: float = 1e-8, weight_decay: float = 0.0, amsgrad: bool = False) -> None:
    """The file `torch::save(adam, path)` writes for a torch::optim::Adam over ONE parameter (the reference builds one
    Adam per parameter group, src/scene/gaussian_model.cpp:323-345): state/<key>/{step, exp_avg, exp_avg_sq} and one
    param group whose options carry lr / betas / eps / weight_decay / amsgrad.  <key> is the string libtorch derives
    from the parameter's address; on load it is matched by POSITION, so any unique string works."""
    C = torch._C
    new = _new_module_factory()
    sub = lambda parent, name, child: parent._register_attribute(name, child._type(), child)
    key = "0"
    top, state, pstate, groups, group, opts = new(), new(), new(), new(), new(), new()
    top._register_attribute("pytorch_version", C.StringType.get(), "1.5.0")
    pstate._register_attribute("step", C.IntType.get(), int(step))
    pstate._register_attribute("exp_avg", C.TensorType.get(), exp_avg.detach().cpu().contiguous())
    pstate._register_attribute("exp_avg_sq", C.TensorType.get(), exp_avg_sq.detach().cpu().contiguous())
    sub(state, key, pstate)
    sub(top, "state", state)
    groups._register_attribute("param_groups/size", C.TensorType.get(), torch.tensor(1))
    group._register_attribute("params/size", C.TensorType.get(), torch.tensor(1))
    group._register_attribute("params/0", C.StringType.get(), key)
    opts._register_attribute("lr", C.FloatType.get(), float(lr))
    opts._register_attribute("betas", C.TupleType([C.FloatType.get(), C.FloatType.get()]), (float(betas[0]), float(betas[1])))
    opts._register_attribute("eps", C.FloatType.get(), float(eps))
    opts._register_attribute("weight_decay", C.FloatType.get(), float(weight_decay))
    opts._register_attribute("amsgrad", C.BoolType.get(), bool(amsgrad))
    sub(group, "options", opts)
    sub(groups, "param_groups/0", group)
    sub(top, "param_groups", groups)
    _save_module(top, path)


def read_adam_archive(path: str):
    """(step, lr, exp_avg, exp_avg_sq) of a one-parameter Adam archive (written here or by libtorch's torch::save)."""
    m = _jit_load(path)
    try:
        group = getattr(m.param_groups, "param_groups/0")
        key = getattr(group, "params/0")
        st = getattr(m.state, key)
        return int(st.step), float(group.options.lr), st.exp_avg.detach(), st.exp_avg_sq.detach()
    except AttributeError as e:
        raise ValueError(f"{path}: not a one-parameter torch::optim::Adam archive ({e})") from None


def capture(model, tensors_path: str, opt_path_for_xyz: str | None = None, opt_path_for_f_dc: str | None = None,
            opt_path_for_f_rest: str | None = None, opt_path_for_opacity: str | None = None,
            opt_path_for_scaling: str | None = None, opt_path_for_rotation: str | None = None,
            fmt: str = "libtorch", moments=None) -> None:
    """CoreParams::capture (src/scene/gaussian_model.cpp:76-132): same arguments, same entry order.
    fmt: "libtorch" (the reference's torch::save archives) or "flat".  moments: (exp_avg[6], exp_avg_sq[6]) to write
    instead of the model's own (the peer-sharded optimiser gathers its shards first)."""
    if fmt not in ("libtorch", "flat"):
        raise ValueError(f"unknown checkpoint format {fmt!r}")
    if fmt == "libtorch":
        write_tensor_archive(tensors_path, [
            torch.tensor(int(model.active_sh_degree), dtype=torch.int32), model.xyz_, model.features_dc_,
            model.features_rest_, model.scaling_, model.rotation_, model.opacity_, model.max_radii2D_,
            model.xyz_gradient_accum_, model.denom_, torch.tensor(float(model.spatial_lr_scale), dtype=torch.float32)])
    else:
        write_entries(tensors_path, [
            np.array(model.active_sh_degree, np.int64), _np(model.xyz_), _np(model.features_dc_),
            _np(model.features_rest_), _np(model.scaling_), _np(model.rotation_), _np(model.opacity_),
            _np(model.max_radii2D_), _np(model.xyz_gradient_accum_), _np(model.denom_),
            np.array(model.spatial_lr_scale, np.float32)])
    paths = (opt_path_for_xyz, opt_path_for_f_dc, opt_path_for_f_rest, opt_path_for_opacity, opt_path_for_scaling,
             opt_path_for_rotation)
    if moments is None and getattr(model, "_opt_ready", False) and model.exp_avg is not None:
        moments = (model.exp_avg, model.exp_avg_sq)
    if moments is not None:
        for k, path in enumerate(paths):
            if not path:
                continue
            m, v = moments[0][k].reshape(model.params()[k].shape), moments[1][k].reshape(model.params()[k].shape)
            if fmt == "libtorch":
                write_adam_archive(path, model.step_count[k], model.lr[k], m, v)
            else:
                write_entries(path, [np.array(model.step_count[k], np.int64), np.array(model.lr[k], np.float64),
                                     _np(m), _np(v)])


def restore(model, tensors_path: str, opt_path_for_xyz: str | None = None, opt_path_for_f_dc: str | None = None,
            opt_path_for_f_rest: str | None = None, opt_path_for_opacity: str | None = None,
            opt_path_for_scaling: str | None = None, opt_path_for_rotation: str | None = None,
            opt_params=None) -> None:
    """CoreParams::restore (src/scene/gaussian_model.cpp:152-202): every file is optional, like there; the container
    of each file (libtorch archive or flat) is recognised from its first bytes.
    opt_params: OptimizationParams -- GaussianModel::restore (gaussian_model.cpp:247-262) calls setup(params) before it
    restores; pass them to get that behaviour.  Without them the optimiser must have been set up already for its
    files to be applied (the reference's optimizers_ map must hold the six Adam objects before torch::load)."""
    dev = model.device
    if opt_params is not None and hasattr(model, "setup"):
        model.setup(opt_params)
    if os.path.exists(tensors_path):
        if _is_flat(tensors_path):
            e = read_entries(tensors_path)
        else:
            e = [x.numpy() for x in read_tensor_archive(tensors_path)]
        if len(e) != 11:
            raise ValueError(f"{tensors_path}: expected 11 entries, found {len(e)}")
        t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
        model.active_sh_degree = int(e[0])
        model.xyz_, model.features_dc_, model.features_rest_ = t(e[1]), t(e[2]), t(e[3])
        model.scaling_, model.rotation_, model.opacity_ = t(e[4]), t(e[5]), t(e[6])
        model.max_radii2D_, model.xyz_gradient_accum_, model.denom_ = t(e[7]).to(torch.int32), t(e[8]), t(e[9])
        model.spatial_lr_scale = float(e[10])
        model.max_sh_degree = int(round((1 + e[3].shape[1]) ** 0.5)) - 1
        if getattr(model, "_opt_ready", False):     # N may differ from the set-up model: re-shape the state
            model.widths = [int(p.numel() // max(model.N, 1)) for p in model.params()]
            model.exp_avg = [torch.zeros_like(p) for p in model.params()]
            model.exp_avg_sq = [torch.zeros_like(p) for p in model.params()]
            model.grads = [torch.zeros_like(p) for p in model.params()]
            model.grad_means2D = torch.zeros((model.N, 2), device=dev)
    paths = (opt_path_for_xyz, opt_path_for_f_dc, opt_path_for_f_rest, opt_path_for_opacity, opt_path_for_scaling,
             opt_path_for_rotation)
    for k, path in enumerate(paths):
        if path and os.path.exists(path):
            if not getattr(model, "_opt_ready", False):
                raise RuntimeError("restore: call GaussianModel.setup() before restoring optimiser state")
            if _is_flat(path):
                step, lr, m, v = read_entries(path)
            else:
                step, lr, m, v = read_adam_archive(path)
                m, v = m.numpy(), v.numpy()
            if tuple(m.shape) != tuple(model.params()[k].shape):
                raise ValueError(f"{path}: moment shape {m.shape} does not match parameter {GROUPS[k]}")
            model.step_count[k] = int(step)
            model.lr[k] = float(lr)
            model.exp_avg[k] = torch.from_numpy(np.ascontiguousarray(m)).to(dev)
            model.exp_avg_sq[k] = torch.from_numpy(np.ascontiguousarray(v)).to(dev)


def read_native(path: str) -> dict:
    """The single-file checkpoint of the native trainer (gs_trainer_save, csrc/trainer.cu) as a dict of numpy arrays."""
    e = read_entries(path)
    if len(e) != 24:
        raise ValueError(f"{path}: not a gs_trainer checkpoint ({len(e)} entries)")
    head = e[0]
    return dict(iteration=int(head[0]), active_sh_degree=int(head[1]), sh_degree_at_0=int(head[2]), N=int(head[3]),
                K=int(head[4]), draws=int(head[5]), pair_capacity=int(head[6]), step_count=[int(x) for x in e[1]],
                params=e[2:8], exp_avg=e[8:14], exp_avg_sq=e[14:20], max_radii2D=e[20].reshape(-1),
                xyz_gradient_accum=e[21], denom=e[22], stack=e[23])


def native_to_libtorch(native_path: str, folder: str, lr: Sequence[float] | None = None, spatial_lr_scale: float = 1.0):
    """Convert a native-trainer checkpoint into the seven torch::save archives of the reference
    (CoreParams::capture's files; names *_<iteration>).  lr: the six learning rates to record in the Adam options
    (default: the reference's, src/arguments/params.h:52-60).  Returns the seven paths."""
    d = read_native(native_path)
    N, K = d["N"], d["K"]
    lr = list(lr) if lr is not None else [1.6e-4 * spatial_lr_scale, 2.5e-3, 2.5e-3 / 20, 0.05, 5e-3, 1e-3]
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a))
    shapes = [(N, 3), (N, 1, 3), (N, K - 1, 3), (N, 1), (N, 3), (N, 4)]
    P = [t(a).reshape(s) for a, s in zip(d["params"], shapes)]
    paths = checkpoint_paths(folder, d["iteration"])
    write_tensor_archive(paths[0], [
        torch.tensor(d["active_sh_degree"], dtype=torch.int32), P[0], P[1], P[2], P[4], P[5], P[3],
        t(d["max_radii2D"]), t(d["xyz_gradient_accum"]).reshape(N, 1), t(d["denom"]).reshape(N, 1),
        torch.tensor(float(spatial_lr_scale), dtype=torch.float32)])
    for k in range(6):
        write_adam_archive(paths[1 + k], d["step_count"][k], lr[k], t(d["exp_avg"][k]).reshape(shapes[k]),
                           t(d["exp_avg_sq"][k]).reshape(shapes[k]))
    return paths


def _is_flat(path: str) -> bool:
    with open(path, "rb") as f:
        return f.read(8) == MAGIC


def checkpoint_paths(folder: str, iteration: int):
    """The seven files of one checkpoint; names end in _<iteration> so that search_for_max_iteration finds them."""
    os.makedirs(folder, exist_ok=True)
    return [os.path.join(folder, f"{name}_{iteration}") for name in ("tensors",) + tuple("opt_" + g for g in GROUPS)]


def search_for_max_iteration(folder: str) -> int:
    """src/utils/system_utils.cpp:7-31: the largest integer after the last '_' of any file name in `folder`;
    -1 for none; a name whose suffix is not a number is an error, as in the reference."""
    max_iter = -1
    for name in os.listdir(folder):
        pos = name.rfind("_")
        if pos != -1:
            tail = name[pos + 1:]
            digits = tail[1:] if tail[:1] in ("+", "-") else tail      # std::stoi: at most ONE sign character
            n = 0
            while n < len(digits) and digits[n].isdigit():   # std::stoi: leading integer, trailing text ignored
                n += 1
            if n == 0:
                raise RuntimeError("Invalid iteration number in " + name)
            max_iter = max(max_iter, int(tail[:len(tail) - len(digits) + n]))
    return max_iter


def save_ply(model, path: str) -> None:
    """Binary little-endian PLY in the attribute order of the published trainer's save_ply."""
    xyz = _np(model.xyz_)
    N = xyz.shape[0]
    f_dc = _np(model.features_dc_).transpose(0, 2, 1).reshape(N, -1)
    f_rest = _np(model.features_rest_).transpose(0, 2, 1).reshape(N, -1)
    cols = [xyz, np.zeros_like(xyz), f_dc, f_rest, _np(model.opacity_).reshape(N, 1), _np(model.scaling_),
            _np(model.rotation_)]
    names = ["x", "y", "z", "nx", "ny", "nz"] + [f"f_dc_{i}" for i in range(f_dc.shape[1])] + \
            [f"f_rest_{i}" for i in range(f_rest.shape[1])] + ["opacity"] + [f"scale_{i}" for i in range(3)] + \
            [f"rot_{i}" for i in range(4)]
    data = np.ascontiguousarray(np.concatenate(cols, axis=1), dtype="<f4")
    header = "ply\nformat binary_little_endian 1.0\n" + f"element vertex {N}\n" + \
             "".join(f"property float {n}\n" for n in names) + "end_header\n"
    with open(path + ".tmp", "wb") as f:
        f.write(header.encode("ascii"))
        f.write(data.tobytes())
    os.replace(path + ".tmp", path)


def load_ply(model, path: str) -> None:
    """Inverse of save_ply (binary little-endian float properties only)."""
    with open(path, "rb") as f:
        names, N = [], 0
        while True:
            line = f.readline().decode("ascii").strip()
            if line.startswith("element vertex"):
                N = int(line.split()[-1])
            elif line.startswith("property float"):
                names.append(line.split()[-1])
            elif line == "end_header":
                break
        data = np.frombuffer(f.read(N * len(names) * 4), dtype="<f4").reshape(N, len(names))
    col = {n: i for i, n in enumerate(names)}
    pick = lambda prefix: data[:, [col[n] for n in sorted((n for n in names if n.startswith(prefix)),
                                                           key=lambda s: int(s.rsplit("_", 1)[1]))]]
    dev = model.device
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a, dtype=np.float32)).to(dev)
    model.xyz_ = t(data[:, [col["x"], col["y"], col["z"]]])
    model.features_dc_ = t(pick("f_dc_").reshape(N, 3, -1).transpose(0, 2, 1))
    rest = pick("f_rest_")
    model.features_rest_ = t(rest.reshape(N, 3, -1).transpose(0, 2, 1)) if rest.shape[1] else torch.empty(N, 0, 3, device=dev)
    model.opacity_ = t(data[:, [col["opacity"]]])
    model.scaling_ = t(pick("scale_"))
    model.rotation_ = t(pick("rot_"))
    model.max_sh_degree = int(round((1 + model.features_rest_.shape[1]) ** 0.5)) - 1
    model.active_sh_degree = model.max_sh_degree
    model.max_radii2D_ = torch.zeros(N, dtype=torch.int32, device=dev)
