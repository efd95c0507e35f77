"""Builds tests/shim/shim_module.cpp (the libtorch shim test extension) in-tree, next to this file."""
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))


def build(verbose=False):
    from torch.utils import cpp_extension
    os.environ.setdefault("CC", "/usr/bin/gcc")
    os.environ.setdefault("CXX", "/usr/bin/g++")
    out = os.path.join(HERE, "_build")
    os.makedirs(out, exist_ok=True)
    libdir = os.path.join(ROOT, "3d_gaussian_splatting_b200")
    return cpp_extension.load(
        name="gsplat_shim_test", sources=[os.path.join(HERE, "shim_module.cpp")], build_directory=out,
        extra_include_paths=[os.path.join(ROOT, "include"), "/usr/local/cuda/include"], extra_cflags=["-O1", "-std=c++17"],
        extra_ldflags=[f"-L{libdir}", "-l:libgsplat_b200.so", f"-Wl,-rpath,{libdir}", "-lc10_cuda",
                       f"-L{os.path.join(os.path.dirname(cpp_extension.__file__), '..', 'lib')}"],
        with_cuda=False, verbose=verbose)


if __name__ == "__main__":
    m = build(verbose="-v" in sys.argv)
    print(m.__file__)
