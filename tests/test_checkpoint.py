"""Checkpoint / export formats (3d_gaussian_splatting_b200/checkpoint.py) on CPU tensors: the capture / restore
pair mirrors GaussianModel::CoreParams (src/scene/gaussian_model.cpp:76-202), search_for_max_iteration mirrors
src/utils/system_utils.cpp:7-31 (its unit test expects 7000 from a point_cloud folder whose newest entry is
iteration_7000)."""
import importlib
import os

import numpy as np
import pytest

torch = pytest.importorskip("torch")
ck = importlib.import_module("3d_gaussian_splatting_b200.checkpoint")


class Model:
    """The fields of GaussianModel that a checkpoint holds (src/scene/gaussian_model.h:10-25) on CPU tensors."""

    def __init__(self, N=0, K=16, seed=0, opt=True):
        g = torch.Generator().manual_seed(seed)
        r = lambda *s: torch.randn(*s, generator=g)
        self.device = torch.device("cpu")
        self.active_sh_degree, self.max_sh_degree, self.spatial_lr_scale = 2, 3, 1.25
        self.xyz_, self.features_dc_, self.features_rest_ = r(N, 3), r(N, 1, 3), r(N, K - 1, 3)
        self.scaling_, self.rotation_, self.opacity_ = r(N, 3), r(N, 4), r(N, 1)
        self.max_radii2D_ = torch.randint(0, 50, (N,), generator=g, dtype=torch.int32)
        self.xyz_gradient_accum_, self.denom_ = r(N, 1).abs(), torch.randint(0, 9, (N, 1), generator=g).float()
        self._opt_ready = opt
        if opt:
            self.exp_avg = [r(*p.shape) for p in self.params()]
            self.exp_avg_sq = [r(*p.shape).abs() for p in self.params()]
            self.step_count = [11, 12, 13, 14, 15, 16]
            self.lr = [1.6e-4, 2.5e-3, 1.25e-4, 5e-2, 5e-3, 1e-3]

    @property
    def N(self):
        return int(self.xyz_.shape[0])

    def params(self):
        return [self.xyz_, self.features_dc_, self.features_rest_, self.opacity_, self.scaling_, self.rotation_]


FIELDS = ["xyz_", "features_dc_", "features_rest_", "scaling_", "rotation_", "opacity_", "max_radii2D_",
          "xyz_gradient_accum_", "denom_"]


@pytest.mark.parametrize("fmt", ["libtorch", "flat"])
def test_capture_restore_roundtrip(tmp_path, fmt):
    a = Model(N=1234, seed=1)
    paths = ck.checkpoint_paths(str(tmp_path), 7000)
    assert len(paths) == 7
    ck.capture(a, *paths, fmt=fmt)
    b = Model(N=10, seed=2)                 # different N: restore re-shapes the optimiser state
    ck.restore(b, *paths)
    for f in FIELDS:
        assert torch.equal(getattr(a, f), getattr(b, f)), f
    assert b.active_sh_degree == 2 and b.spatial_lr_scale == 1.25 and b.max_sh_degree == 3
    for k in range(6):
        assert torch.equal(a.exp_avg[k], b.exp_avg[k]) and torch.equal(a.exp_avg_sq[k], b.exp_avg_sq[k])
    assert b.step_count == a.step_count and b.lr == a.lr
    # entry order of the tensors file == the reference's vector (gaussian_model.cpp:85-97)
    e = ck.read_entries(paths[0]) if fmt == "flat" else [x.numpy() for x in ck.read_tensor_archive(paths[0])]
    assert len(e) == 11 and int(e[0]) == 2 and e[1].shape == (1234, 3) and e[3].shape == (1234, 15, 3)
    assert e[4].shape == (1234, 3) and e[5].shape == (1234, 4) and e[6].shape == (1234, 1) and e[7].dtype == np.int32
    assert float(e[10]) == 1.25


def test_restore_tolerates_missing_files_like_the_reference(tmp_path):
    a = Model(N=50, seed=3)
    paths = ck.checkpoint_paths(str(tmp_path), 100)
    ck.capture(a, paths[0], paths[1])        # tensors + the xyz optimiser only
    b = Model(N=50, seed=4)
    before = [m.clone() for m in b.exp_avg]
    ck.restore(b, *paths)
    assert torch.equal(b.xyz_, a.xyz_) and torch.equal(b.exp_avg[0], a.exp_avg[0])
    assert all(not m.any() for m in b.exp_avg[1:])       # re-shaped to zeros, no file to fill them
    c = Model(N=50, seed=5)
    ck.restore(c, str(tmp_path / "nothing"), *[str(tmp_path / "none")] * 6)   # nothing exists: nothing changes
    assert torch.equal(c.xyz_, Model(N=50, seed=5).xyz_)


def test_empty_model_and_corruption(tmp_path):
    a = Model(N=0, seed=1)
    p = ck.checkpoint_paths(str(tmp_path), 1)
    ck.capture(a, *p, fmt="flat")
    b = Model(N=3, seed=2)
    ck.restore(b, *p)
    assert b.N == 0
    with open(p[0], "r+b") as f:
        f.write(b"garbage!")
    with pytest.raises(ValueError):
        ck.restore(b, *p)
    blob = open(p[1], "rb").read()
    open(p[1], "wb").write(blob[:-1] if len(blob) > 40 else blob[:20])
    ck.capture(Model(N=5, seed=1), *ck.checkpoint_paths(str(tmp_path / "t"), 2), fmt="flat")
    q = ck.checkpoint_paths(str(tmp_path / "t"), 2)
    blob = open(q[1], "rb").read()
    open(q[1], "wb").write(blob[:-4])
    with pytest.raises(ValueError):
        ck.restore(Model(N=5, seed=1), *q)


def test_search_for_max_iteration(tmp_path):
    d = tmp_path / "point_cloud"
    d.mkdir()
    assert ck.search_for_max_iteration(str(d)) == -1
    for it in (7000, 30, 1000):
        (d / f"iteration_{it}").mkdir()
    (d / "noseparator").write_text("")
    assert ck.search_for_max_iteration(str(d)) == 7000      # the value the reference's unit test expects
    (d / "iteration_12abc").write_text("")                   # std::stoi reads the leading integer
    assert ck.search_for_max_iteration(str(d)) == 7000
    (d / "iteration_final").write_text("")
    with pytest.raises(RuntimeError):
        ck.search_for_max_iteration(str(d))


def test_ply_roundtrip(tmp_path):
    a = Model(N=321, seed=9, opt=False)
    path = str(tmp_path / "point_cloud.ply")
    ck.save_ply(a, path)
    head = open(path, "rb").read(2000).decode("ascii", "replace")
    assert head.startswith("ply\nformat binary_little_endian 1.0\nelement vertex 321\n")
    assert os.path.getsize(path) == head.index("end_header\n") + len("end_header\n") + 321 * 62 * 4
    b = Model(N=1, seed=1, opt=False)
    ck.load_ply(b, path)
    for f in ["xyz_", "features_dc_", "features_rest_", "scaling_", "rotation_", "opacity_"]:
        assert torch.equal(getattr(a, f), getattr(b, f)), f
    # f_rest is stored channel-major, as the published exporter does (transpose(1, 2) before flattening)
    raw = np.frombuffer(open(path, "rb").read()[-321 * 62 * 4:], dtype="<f4").reshape(321, 62)
    np.testing.assert_array_equal(raw[:, 9:9 + 15], a.features_rest_[:, :, 0].numpy())


def _libtorch():
    """tests/ckpt/ckpt_module.cpp: libtorch's own torch::save / torch::load, the calls the reference makes."""
    import sys
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "ckpt"))
    import build_ckpt
    return build_ckpt.build()


def test_checkpoints_are_libtorch_archives_both_ways(tmp_path):
    """The default container IS the reference's: the tensors file written here loads with
    torch::load(std::vector<torch::Tensor>&) into the 11 entries of CoreParams::restore (gaussian_model.cpp:163-176),
    each optimiser file loads into a torch::optim::Adam set up as GaussianModel::setup does (:323-345) -- and files
    written by libtorch's torch::save restore into this model."""
    ext = _libtorch()
    a = Model(N=321, seed=6)
    paths = ck.checkpoint_paths(str(tmp_path), 30000)
    ck.capture(a, *paths)
    t = ext.load_tensors(paths[0])
    assert len(t) == 11 and int(t[0]) == a.active_sh_degree and abs(float(t[10]) - a.spatial_lr_scale) < 1e-7
    order = [a.xyz_, a.features_dc_, a.features_rest_, a.scaling_, a.rotation_, a.opacity_, a.max_radii2D_,
             a.xyz_gradient_accum_, a.denom_]
    for got, want in zip(t[1:10], order):
        assert got.dtype == want.dtype and torch.equal(got, want)
    for k, p in enumerate(a.params()):
        step, lr, m, v = ext.load_adam(paths[1 + k], p.clone(), 0.5)      # lr 0.5: must be overwritten by the file's
        assert int(step) == a.step_count[k] and abs(float(lr) - a.lr[k]) < 1e-15
        assert torch.equal(m, a.exp_avg[k]) and torch.equal(v, a.exp_avg_sq[k])
    # the other direction: libtorch writes, this package restores
    q = [str(tmp_path / f"cpp_{i}") for i in range(7)]
    ext.save_tensors(q[0], [torch.tensor(a.active_sh_degree, dtype=torch.int32)] + order +
                     [torch.tensor(a.spatial_lr_scale, dtype=torch.float32)])
    for k, p in enumerate(a.params()):
        ext.save_adam(q[1 + k], p.clone(), a.lr[k], a.step_count[k], a.exp_avg[k], a.exp_avg_sq[k])
    b = Model(N=7, seed=8)
    ck.restore(b, *q)
    for f in FIELDS:
        assert torch.equal(getattr(a, f), getattr(b, f)), f
    assert b.active_sh_degree == a.active_sh_degree and b.step_count == a.step_count
    assert all(abs(x - y) < 1e-15 for x, y in zip(b.lr, a.lr))
    for k in range(6):
        assert torch.equal(a.exp_avg[k], b.exp_avg[k]) and torch.equal(a.exp_avg_sq[k], b.exp_avg_sq[k])


def test_search_for_max_iteration_sign_handling(tmp_path):
    """std::stoi accepts ONE sign character (src/utils/system_utils.cpp:7-31 parses the suffix with it)."""
    d = tmp_path / "pc"
    d.mkdir()
    (d / "iteration_+15").write_text("")
    assert ck.search_for_max_iteration(str(d)) == 15
    (d / "iteration_+-5").write_text("")
    with pytest.raises(RuntimeError):
        ck.search_for_max_iteration(str(d))
