"""In-tree build of libgsplat_b200.so (sm_100a only).

`python -m 3d_gaussian_splatting_b200.build` or build_library() compiles every .cu/.cpp under
csrc/ with nvcc for sm_100a (-lineinfo so ncu source pages map to our code) and links one
shared library next to this file.  The built .so is git-ignored but travels to the GPU box.
preprocess.cu is compiled with --fmad=false: its integer outputs (radii, tile counts, sort
keys) must be bit-identical to the CPU oracle, so no fused multiply-add contraction there;
preprocess_bwd.cu likewise, so that its three kernel variants agree bit for bit.
"""
from __future__ import annotations

import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OBJ = os.path.join(HERE, "_obj")
LIB = os.path.join(HERE, "libgsplat_b200.so")

ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
COMMON = ["-O3", "-std=c++17", "-lineinfo", "-Xcompiler", "-fPIC,-fvisibility=hidden", "--use_fast_math=false"]
PER_FILE = {
    "preprocess.cu": ["--fmad=false"],
    # K4b exists in three kernels (generic, TMA-staged, fused with Adam) that must produce the same gradient bits:
    # without contraction every operation is a single IEEE operation, whatever the surrounding code
    "preprocess_bwd.cu": ["--fmad=false"],
}
# host compiler: the image exports CC/CXX pointing at a gcc without libgomp etc.; use the system one
CCBIN = "/usr/bin/g++"


def _sources():
    return sorted(f for f in os.listdir(CSRC) if f.endswith((".cu", ".cpp")))


def _stale(out: str, deps) -> bool:
    if not os.path.exists(out):
        return True
    t = os.path.getmtime(out)
    return any(os.path.getmtime(d) > t for d in deps)


def build_library(force: bool = False, verbose: bool = False) -> str:
    os.makedirs(OBJ, exist_ok=True)
    headers = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cuh", ".h"))]
    headers.append(os.path.join(HERE, "..", "include", "gsplat_b200.h"))
    headers.append(os.path.abspath(__file__))
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    jobs = []
    objs = []
    for src in _sources():
        obj = os.path.join(OBJ, src.rsplit(".", 1)[0] + ".o")
        objs.append(obj)
        if force or _stale(obj, [os.path.join(CSRC, src)] + headers):
            cmd = [nvcc, "-ccbin", CCBIN, *ARCH, *[c for c in COMMON if c != "--use_fast_math=false"],
                   *PER_FILE.get(src, []), "-c", os.path.join(CSRC, src), "-o", obj]
            if verbose:
                cmd.insert(1, "-Xptxas=-v")
            jobs.append(cmd)

    def run(cmd):
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("nvcc failed:\n" + " ".join(cmd) + "\n" + r.stdout + r.stderr)
        return r.stderr

    if jobs:
        with ThreadPoolExecutor(max_workers=min(8, len(jobs))) as ex:
            for log in ex.map(run, jobs):
                if verbose and log:
                    print(log, file=sys.stderr)
    if jobs or not os.path.exists(LIB):
        run([nvcc, "-ccbin", CCBIN, *ARCH, "-shared", "-o", LIB, *objs, "-cudart", "static"])
    return LIB


if __name__ == "__main__":
    print(build_library(force="--force" in sys.argv, verbose="-v" in sys.argv))
