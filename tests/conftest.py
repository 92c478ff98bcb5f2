import importlib
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests", "emu"), os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)

GOLDEN = os.path.join(ROOT, "tests", "golden")
pkg = importlib.import_module("cpu-ray-tracing-implementation_b200")
abi = pkg._abi
Scene = pkg.scene.Scene

ALL_SCENES = sorted(f[:-4] for f in os.listdir(GOLDEN) if f.endswith(".npz"))
CONFIG_SCENES = ["cornell", "rtow", "rtnw", "fox"]  # BASELINE.json configs[0..3]; configs[4] (Sponza) is built procedurally
VOLUME_SCENES = {"rtnw", "cornell_volume", "zoo_nolight", "zoo_quad", "zoo_sphere"}


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200); run with -m gpu")


class Golden:
    """One tests/golden/<name>.npz: flattened scene + cameras + outputs of the reference itself."""

    def __init__(self, name):
        self.name = name
        self.z = np.load(os.path.join(GOLDEN, name + ".npz"))
        self.scene = Scene.load(self.z)
        self.scene.name = name
        self.cam_small = abi.Camera.from_bytes(self.z["cam_small"].tobytes())
        self.cam_full = abi.Camera.from_bytes(self.z["cam_full"].tobytes())

    def __getitem__(self, k):
        return self.z[k]

    def volume_prims(self):
        n = self.scene.nodes
        return set(int(c) for c in n["c"][n["kind"] == abi.NODE_VOLUME])


_cache = {}


def golden(name):
    if name not in _cache:
        _cache[name] = Golden(name)
    return _cache[name]


@pytest.fixture(scope="session")
def build_oracle():
    import subprocess

    subprocess.run(["make", "-s", "-f", os.path.join(ROOT, "oracle", "Makefile")], check=True, cwd=ROOT)
    import oracleport

    return oracleport


def same(a, b):
    """Bit-equality that treats NaN == NaN."""
    a, b = np.asarray(a), np.asarray(b)
    if a.dtype.kind == "f":
        return (a == b) | (np.isnan(a) & np.isnan(b))
    return a == b


def z_scores(sum_a, sumsq_a, n_a, sum_b, sumsq_b, n_b):
    """Per-channel z of mean_a - mean_b.  The per-pixel variance is pooled over both renders: radiance samples are
    heavy-tailed (rare light-sampled spikes), and a pixel whose spike landed on one side only would otherwise be
    judged against a variance that has never seen one."""
    mean_a, mean_b = sum_a / n_a, sum_b / n_b
    var_a = np.maximum(sumsq_a / n_a - mean_a**2, 0.0)
    var_b = np.maximum(sumsq_b / n_b - mean_b**2, 0.0)
    var = (n_a * var_a + n_b * var_b) / (n_a + n_b)
    se = np.sqrt(var / n_a + var / n_b)
    with np.errstate(divide="ignore", invalid="ignore"):
        z = np.where(se > 0, (mean_a - mean_b) / se, 0.0)
    return z, mean_a, mean_b


def map_scatter_draws(scene, rec12, mode, draws, counts, quad_random_swapped=True):
    """The std::rand values one level of ray_color consumed (reference consumption order, as logged by
    oracle/ref_harness.cc: ref_scatter_logged) -> the draws6 = {m0 m1 m2 s0 s1 s2} addressing of b200rt_scatter.

    Consumption order in the reference: material::scatter first (metal: u1, u2 of random_in_unit_sphere, src/material.h:87;
    dielectric: one value unless total internal reflection short-circuits it, :123; gloss: the specular test, then r1, r2 of
    random_cosine_direction when specular, :161-163), then for kRandom the pdf: dual_pdf's choice when a light was passed
    (src/pdf.h:53), then the pair of generate().  quad::random writes `random_double() * u_ + random_double() * v_` in ONE
    expression (src/quad.h:76): the order of the two calls is the compiler's; g++ 13 -O2 calls the one multiplying v_ first
    (established by test_scatter_draw_order_of_quad_random)."""
    n = len(rec12)
    out = np.zeros((n, 6))
    kinds = scene.materials["kind"][rec12[:, 10].astype(int)]
    has_light = scene.light_kind != abi.LIGHT_NONE
    for i in range(n):
        if mode[i] < 0:
            continue
        log = list(draws[i, : counts[i]])
        k = kinds[i]
        if k == abi.MAT_METAL:
            out[i, 1:3] = log[0:2]
            log = log[2:]
        elif k == abi.MAT_DIELECTRIC:
            if log:
                out[i, 0] = log[0]
            else:
                out[i, 0] = 0.5  # never read: total internal reflection decided it
            log = log[1:]
        elif k == abi.MAT_GLOSS:
            out[i, 0] = log[0]
            log = log[1:]
            if mode[i] == 1:
                out[i, 1:3] = log[0:2]
                log = log[2:]
        if mode[i] == 2:
            if has_light:
                out[i, 3] = log[0]
                log = log[1:]
                pair = log[0:2]
                if out[i, 3] < 0.5 and scene.light_kind == abi.LIGHT_QUAD and quad_random_swapped:
                    pair = pair[::-1]
            else:
                pair = log[0:2]
            out[i, 4:6] = pair
            log = log[2:]
        assert not log, (i, k, mode[i], counts[i])
    return out


def scatter_keep(g):
    """Records of the golden scatter batch that can be compared: the ray hit something, and - for value_noise textures, whose
    lattice lookup has no bounds check (src/noise.h:109-116: points on the last lattice cell read past the array, in the
    reference and in any faithful restatement) - the hit point lies inside the lattice."""
    ref, rec = g["scatter_out16"], g["scatter_rec12"]
    keep = ref[:, 0] >= 0
    if len(g.scene.value_noise):
        res = int(g.scene.value_noise["resolution"].min())
        keep &= ((rec[:, 0:3] >= 0) & (rec[:, 0:3] < res - 2)).all(axis=1)
    return keep
