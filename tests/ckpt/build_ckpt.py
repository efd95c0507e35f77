"""Builds tests/ckpt/ckpt_module.cpp (libtorch's own torch::save / torch::load, CPU only) in-tree."""
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))


def build(verbose=False):
    from torch.utils import cpp_extension
    os.environ.setdefault("CC", "/usr/bin/gcc")
    os.environ.setdefault("CXX", "/usr/bin/g++")
    out = os.path.join(HERE, "_build")
    os.makedirs(out, exist_ok=True)
    return cpp_extension.load(name="gsplat_ckpt_test", sources=[os.path.join(HERE, "ckpt_module.cpp")],
                              build_directory=out, extra_cflags=["-O1", "-std=c++17"], with_cuda=False, verbose=verbose)


if __name__ == "__main__":
    print(build(verbose="-v" in sys.argv).__file__)
