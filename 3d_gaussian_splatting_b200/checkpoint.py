"""Checkpoint / export formats next to the parameter store.

capture() / restore() mirror GaussianModel::CoreParams::capture / restore
(src/scene/gaussian_model.cpp:76-202): ONE tensors file holding the same eleven entries in the same order
(active_sh_degree, xyz, features_dc, features_rest, scaling, rotation, opacity, max_radii2D,
xyz_gradient_accum, denom, spatial_lr_scale; :85-97) plus SIX optimiser files (xyz, f_dc, f_rest, opacity,
scaling, rotation; :100-131), each restored only if it exists (:163-201).  The reference writes libtorch
archives (torch::save); libtorch's pickle/zip container is not a format worth re-implementing for a drop-in
that owns flat SoA buffers, so the container here is a flat little-endian dump:

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


def capture(model, tensors_path: str, opt_path_for_xyz: str | None = None, opt_path_for_f_dc: str | None = None,
            opt_path_for_f_rest: str | None = None, opt_path_for_opacity: str | None = None,
            opt_path_for_scaling: str | None = None, opt_path_for_rotation: str | None = None) -> None:
    """CoreParams::capture (src/scene/gaussian_model.cpp:76-132): same arguments, same entry order."""
    write_entries(tensors_path, [
        np.array(model.active_sh_degree, np.int64), _np(model.xyz_), _np(model.features_dc_),
        _np(model.features_rest_), _np(model.scaling_), _np(model.rotation_), _np(model.opacity_),
        _np(model.max_radii2D_), _np(model.xyz_gradient_accum_), _np(model.denom_),
        np.array(model.spatial_lr_scale, np.float32)])
    paths = (opt_path_for_xyz, opt_path_for_f_dc, opt_path_for_f_rest, opt_path_for_opacity, opt_path_for_scaling,
             opt_path_for_rotation)
    if getattr(model, "_opt_ready", False) and model.exp_avg is not None:
        for k, path in enumerate(paths):
            if path:
                write_entries(path, [np.array(model.step_count[k], np.int64), np.array(model.lr[k], np.float64),
                                     _np(model.exp_avg[k]), _np(model.exp_avg_sq[k])])


def restore(model, tensors_path: str, opt_path_for_xyz: str | None = None, opt_path_for_f_dc: str | None = None,
            opt_path_for_f_rest: str | None = None, opt_path_for_opacity: str | None = None,
            opt_path_for_scaling: str | None = None, opt_path_for_rotation: str | None = None) -> None:
    """CoreParams::restore (src/scene/gaussian_model.cpp:152-202): every file is optional, like there.  The
    optimiser must have been set up (GaussianModel.setup) for its files to be applied, as in the reference, whose
    optimizers_ map must hold the six Adam objects before torch::load fills them."""
    dev = model.device
    if os.path.exists(tensors_path):
        e = read_entries(tensors_path)
        if len(e) != 11:
            raise ValueError(f"{tensors_path}: expected 11 entries, found {len(e)}")
        t = lambda a: torch.from_numpy(a).to(dev)
        model.active_sh_degree = int(e[0])
        model.xyz_, model.features_dc_, model.features_rest_ = t(e[1]), t(e[2]), t(e[3])
        model.scaling_, model.rotation_, model.opacity_ = t(e[4]), t(e[5]), t(e[6])
        model.max_radii2D_, model.xyz_gradient_accum_, model.denom_ = t(e[7]), t(e[8]), t(e[9])
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
            step, lr, m, v = read_entries(path)
            if tuple(m.shape) != tuple(model.params()[k].shape):
                raise ValueError(f"{path}: moment shape {m.shape} does not match parameter {GROUPS[k]}")
            model.step_count[k] = int(step)
            model.lr[k] = float(lr)
            model.exp_avg[k] = torch.from_numpy(m).to(dev)
            model.exp_avg_sq[k] = torch.from_numpy(v).to(dev)


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
            digits = tail.lstrip("+-")
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
