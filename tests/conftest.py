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
VOLUME_SCENES = {"rtnw", "cornell_volume"}


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
