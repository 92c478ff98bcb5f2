"""CPU execution of the DEVICE FUNCTIONS of csrc/ (tests/emu: the same traverse / intersect / shading / camera source compiled
with g++ behind stand-ins for the CUDA built-ins) against the reference's golden records and the oracle.  No kernel runs here;
this pins the host logic (flattener, BVH builder, fp32 conservative box filter, tie-break by reference order) without a GPU."""
import numpy as np
import pytest
from conftest import ALL_SCENES, VOLUME_SCENES, golden, map_scatter_draws, same, scatter_keep

import emu


@pytest.fixture(scope="module")
def emu_lib():
    emu.build()
    return emu


@pytest.mark.parametrize("leaf", [1, 2, 4])
@pytest.mark.parametrize("name", ALL_SCENES)
def test_emu_hits_vs_reference_golden(emu_lib, name, leaf):
    g = golden(name)
    es = emu_lib.EmuScene(g.scene, leaf_size=leaf)
    out = es.trace(g["trace_rays"], seed=99)
    keep = np.ones(len(out["prim"]), bool)
    if name in VOLUME_SCENES:
        vol = np.array(sorted(g.volume_prims()))
        keep = ~np.isin(out["prim"], vol) & ~np.isin(g["trace_prim"], vol)
    for k in ("prim", "front", "t", "p", "normal", "uv"):
        ok = same(out[k][keep], g["trace_" + k][keep])
        assert ok.all(), f"{name}: {k} differs on {(~ok).sum()} of {keep.sum()} rays"


def test_threaded_host_build_gives_the_serial_tree(emu_lib):
    """bvh_build.h: subtrees of >= 16384 primitives in the top levels are built by their own threads; node numbering, boxes and
    the entry permutation must be those of the one-thread build (262 268 primitives: the Sponza stand-in)."""
    import importlib

    pkg = importlib.import_module("cpu-ray-tracing-implementation_b200")
    sc = pkg.scene.sponza_standin_scene()
    a, b = emu_lib.EmuScene(sc, leaf_size=2), emu_lib.EmuScene(sc, leaf_size=102)
    assert a.info() == b.info()
    assert a.bvh_hash() == b.bvh_hash()


@pytest.mark.parametrize("leaf", [1, 2])
@pytest.mark.parametrize("name", ALL_SCENES)
def test_emu_agglomerative_tree(emu_lib, name, leaf):
    """bvh_build.h: build_agglomerative (what b200rt_upload_scene also builds for scenes of up to 64 primitives and keeps when its
    SAH cost is lower): yet another tree, the same records as the reference."""
    g = golden(name)
    if len(g.scene.nodes) > 4000:
        pytest.skip("O(n^3) builder: small scenes only")
    es = emu_lib.EmuScene(g.scene, leaf_size=-10 - leaf)
    info = es.info()
    assert info["bvh_nodes"] == max(0, info["leaves"] - 1) and 1 <= info["bvh_depth"] <= 64
    out = es.trace(g["trace_rays"], seed=99)
    keep = np.ones(len(out["prim"]), bool)
    if name in VOLUME_SCENES:
        vol = np.array(sorted(g.volume_prims()))
        keep = ~np.isin(out["prim"], vol) & ~np.isin(g["trace_prim"], vol)
    for k in ("prim", "front", "t", "p", "normal", "uv"):
        ok = same(out[k][keep], g["trace_" + k][keep])
        assert ok.all(), f"{name}: {k} differs on {(~ok).sum()} of {keep.sum()} rays"


@pytest.mark.parametrize("leaf", [1, 2, 3])
@pytest.mark.parametrize("name", ALL_SCENES)
def test_emu_clustered_tree_through_finish_tree(emu_lib, name, leaf):
    """The host pass the library runs on the device PLOC topology (csrc/bvh_build.h: finish_tree: leaf merging by the SAH rule,
    depth-first layout, entry permutation) on a serial restatement of the clustering (tests/emu/emu.cc): a different tree again,
    the same records as the reference."""
    g = golden(name)
    es = emu_lib.EmuScene(g.scene, leaf_size=-leaf)
    info, host = es.info(), emu_lib.EmuScene(g.scene, leaf_size=leaf).info()
    assert info["entries"] == host["entries"] and -(-info["entries"] // leaf) <= info["leaves"] <= info["entries"]
    assert info["bvh_nodes"] == max(0, info["leaves"] - 1) and 1 <= info["bvh_depth"] <= 64
    out = es.trace(g["trace_rays"], seed=99)
    keep = np.ones(len(out["prim"]), bool)
    if name in VOLUME_SCENES:
        vol = np.array(sorted(g.volume_prims()))
        keep = ~np.isin(out["prim"], vol) & ~np.isin(g["trace_prim"], vol)
    for k in ("prim", "front", "t", "p", "normal", "uv"):
        ok = same(out[k][keep], g["trace_" + k][keep])
        assert ok.all(), f"{name}: {k} differs on {(~ok).sum()} of {keep.sum()} rays"


@pytest.mark.parametrize("name", ["cornell", "rtow", "fox", "rtnw"])
def test_emu_mirror_equals_fast(emu_lib, name):
    g = golden(name)
    es = emu_lib.EmuScene(g.scene)
    a, b = es.trace(g["trace_rays"], seed=5), es.trace(g["trace_rays"], seed=5, mirror=True)
    for k in ("prim", "t", "front"):
        assert same(a[k], b[k]).all(), k


def test_emu_degenerate_rays_are_conservative(emu_lib, build_oracle):
    """Axis-parallel, zero-component, huge, tiny and NaN rays: the fp32 box filter must never lose a hit the oracle finds."""
    g = golden("rtow")
    es = emu_lib.EmuScene(g.scene)
    rng = np.random.RandomState(7)
    n = 4000
    rays = np.zeros((n, 7))
    rays[:, 0:3] = rng.uniform(-12, 12, size=(n, 3)) * [1, 0.2, 1] + [0, 1.0, 0]
    d = rng.normal(size=(n, 3))
    d[rng.rand(n) < 0.3, 0] = 0.0
    d[rng.rand(n) < 0.3, 2] = 0.0
    d[rng.rand(n) < 0.1, 1] = 0.0
    scale = 10.0 ** rng.uniform(-35, 35, size=(n, 1))
    scale[rng.rand(n) < 0.5] = 1.0
    rays[:, 3:6] = d * scale
    rays[::97, 3] = np.nan
    rays[5::101, 0] = 1e9
    out = es.trace(rays, seed=3)
    ref = build_oracle.trace(g.scene, rays, seed=3)
    assert (out["prim"] == ref["prim"]).all()
    assert same(out["t"], ref["t"]).all()


@pytest.mark.parametrize("name", ["cornell", "rtow", "three_ball_lens", "skybox_fisheye"])
def test_emu_camera_rays(emu_lib, name, build_oracle):
    g = golden(name)
    a = emu_lib.generate_rays(g.cam_small, 11, 0, 2)
    b = build_oracle.generate_rays(g.cam_small, 11, 0, 2)
    assert same(a, b).all()


@pytest.mark.parametrize("name", ["cornell", "rtow", "cornell_volume", "light_earth"])
def test_emu_render_equals_oracle(emu_lib, name, build_oracle):
    g = golden(name)
    a = emu_lib.EmuScene(g.scene).render(g.cam_small, 21, 0, 2)
    b = build_oracle.render(g.scene, g.cam_small, 21, 0, 2, threads=0)
    close = np.abs(a["sum"] - b["sum"]) <= 1e-9 * np.maximum(np.abs(b["sum"]), 1e-3)  # sincos / fused libm calls may differ in the last place
    assert close.mean() > 0.995, close.mean()
    assert abs(a["rays"] - b["stats"]["rays"]) <= 0.002 * b["stats"]["rays"] + 2


@pytest.mark.parametrize("name", ALL_SCENES)
def test_emu_shading_on_reference_draws(emu_lib, name):
    """The DEVICE shading function (csrc/shading.cuh: shade_hit, as k_scatter calls it) executed on the CPU, on the reference's
    own hit records and the std::rand values it consumed: identical to what the reference's material / pdf objects produced
    (here both sides use glibc's libm, so even the sincos / pow paths must agree to the last bit, except where the device code
    deliberately shares work: sincos() instead of separate sin() and cos())."""
    g = golden(name)
    ref, rec = g["scatter_out16"], g["scatter_rec12"]
    keep = scatter_keep(g)
    d6 = map_scatter_draws(g.scene, rec, ref[:, 0], g["scatter_draws"], g["scatter_counts"])
    out = emu_lib.EmuScene(g.scene).scatter(g["scatter_rays"][keep], rec[keep], d6[keep])
    ref = ref[keep]
    assert np.array_equal(out[:, 0], ref[:, 0])
    assert same(out[:, 1:7], ref[:, 1:7]).all(), "emitted / attenuation"
    close = np.isclose(out[:, 7:15], ref[:, 7:15], rtol=1e-13, atol=1e-15) | same(out[:, 7:15], ref[:, 7:15])
    assert close.all(), f"{name}: {(~close).any(axis=1).sum()} records differ"
    assert same(out[:, 7:15], ref[:, 7:15]).all(axis=1).mean() > 0.9
