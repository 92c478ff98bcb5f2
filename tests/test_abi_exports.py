"""-m "not gpu": libb200rt.so loads without a GPU and exports every entry point include/b200rt.h declares (no compute calls);
the ctypes mirror of the POD structures has the sizes the library was compiled with; the product package does not import the oracle."""
import ctypes as C
import os
import re

import pytest
from conftest import ROOT, abi, pkg


def declared_functions():
    text = open(os.path.join(ROOT, "include", "b200rt.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(b200rt_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    lib = pkg.backend.lib()  # raises if the .so is missing: there is no fallback
    names = declared_functions()
    assert len(names) >= 16 and "b200rt_render" in names and "b200rt_format_ppm" in names
    for n in names:
        assert hasattr(lib, n), f"{n} is declared in include/b200rt.h but not exported by libb200rt.so"
    assert set(abi.EXPORTS) == set(names), set(abi.EXPORTS) ^ set(names)


def test_struct_sizes_match_the_compiled_library():
    lib = pkg.backend.lib()
    assert lib.b200rt_abi_version() == abi.ABI_VERSION
    sizes = [C.sizeof(abi.SceneDesc), C.sizeof(abi.Camera), C.sizeof(abi.Stats), abi.node_dtype.itemsize, abi.material_dtype.itemsize,
             abi.texture_dtype.itemsize, abi.image_dtype.itemsize, abi.volume_dtype.itemsize, abi.value_noise_dtype.itemsize, C.sizeof(abi.BvhInfo)]
    for which, size in enumerate(sizes):
        assert lib.b200rt_abi_sizeof(which) == size, which
    assert lib.b200rt_ppm_bound(3, 2) >= len("P3\n3 2\n255\n") + 6 * 36 and lib.b200rt_ppm_bound(0, 5) == 0


def test_no_gpu_means_a_loud_error_not_a_fallback():
    import torch

    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(pkg.backend.B200RTError):
        pkg.backend.Context(0)


def test_product_package_never_touches_the_oracle():
    pkgdir = os.path.join(ROOT, "cpu-ray-tracing-implementation_b200")
    for dirpath, _, files in os.walk(pkgdir):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dirpath, f), errors="replace").read()
                assert "oracleport" not in src and "liboracle" not in src and "refharness" not in src, f
