"""The C-ABI library loads on a CPU-only box, exports every symbol include/gsplat_b200.h declares,
its host-only functions work, and its compute entry points fail loudly without a GPU."""
import ctypes as C
import os
import re
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    src = open(os.path.join(ROOT, "include", "gsplat_b200.h")).read()
    return sorted(set(re.findall(r"GS_API[^;(]*?\b(gs_[a-z0-9_]+)\s*\(", src)))


def test_header_symbols_exported(pkg):
    names = declared_symbols()
    assert len(names) >= 17
    lib = C.CDLL(pkg._lib.LIB_PATH)
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/gsplat_b200.h but not exported"
    assert set(names) == set(pkg._lib.EXPORTS)
    out = subprocess.run(["nm", "-D", "--defined-only", pkg._lib.LIB_PATH], capture_output=True, text=True).stdout
    exported = {l.split()[-1] for l in out.splitlines() if " T " in l}
    assert set(names) <= exported
    # nothing else leaks (hidden visibility): every exported text symbol is part of the ABI
    leaked = {s for s in exported if not s.startswith("gs_")}
    assert not leaked, leaked


def test_struct_layouts_match_header(pkg):
    L = pkg._lib
    assert C.sizeof(L.GsSettings) == 4 * (2 + 2 + 3 + 1 + 16 + 16 + 3 + 2 + 2)
    assert C.sizeof(L.GsGaussians) == 8 * 6 + 4 * 2 + 8 * 2 + 8      # 6 ptrs, 2 ints, 2 ptrs, int + pad
    assert C.sizeof(L.GsGradients) == 8 * 6 + 4 * 2 + 8 * 3 + 8


def test_is_sm100a_only(pkg):
    out = subprocess.run(["cuobjdump", "--list-elf", pkg._lib.LIB_PATH], capture_output=True, text=True).stdout
    archs = set(re.findall(r"sm_\d+a?", out))
    assert archs == {"sm_100a"}, archs


def test_no_cpu_fallback(pkg):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    ctx = C.c_void_p()
    rc = pkg._lib.lib.gs_create(C.byref(ctx), 0)
    assert rc == pkg._lib.GS_ECUDA and not ctx.value
    assert b"no CPU fallback" in pkg._lib.lib.gs_last_error(None)
    with pytest.raises(RuntimeError):
        pkg.Rasterizer(0)


def test_product_does_not_touch_oracle():
    """oracle/ is test infrastructure: nothing under the package may import, include or link it."""
    pk = os.path.join(ROOT, "3d_gaussian_splatting_b200")
    for dirpath, _, files in os.walk(pk):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h")):
                txt = open(os.path.join(dirpath, f), errors="ignore").read()
                assert "oracle_lib" not in txt and "liboracle" not in txt and '#include "orc_' not in txt, f
                assert not re.search(r"^\s*(from|import)\s+oracle\b", txt, re.M), f
