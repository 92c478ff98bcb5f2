"""-m gpu: the CUDA path (libb200rt.so through its C ABI) against the oracle and the reference's golden vectors.

Tolerances: primitive ids, front_face: bit-exact.  t, p, normal, quad u/v: bit-exact (same binary64 operations in the
same order, no FMA).  sphere u/v and fisheye rays go through acos/atan2/asin/sin/cos, where CUDA's libm and glibc may
differ in the last place: 4 ulp.  Radiance per sample: 1e-9 relative on >= 99.5 % of samples (a last-place difference
in a sampled direction can flip a later branch on a few paths); images: RMSE < 1e-3 and mean within 0.5 %.
"""
import numpy as np
import pytest
from conftest import ALL_SCENES, CONFIG_SCENES, VOLUME_SCENES, abi, golden, pkg, same

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx():
    c = pkg.backend.Context(0)
    yield c
    c.close()


def ulp_close(a, b, ulps):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    tol = ulps * np.spacing(np.maximum(np.abs(a), np.abs(b)))
    return (np.abs(a - b) <= tol) | (np.isnan(a) & np.isnan(b))


def check_records(name, out, ref, keep):
    for k in ("prim", "front", "t", "p", "normal"):
        ok = same(out[k][keep], ref[k][keep])
        assert ok.all(), f"{name}: {k} differs on {(~ok).sum()} of {keep.sum()} rays"
    ok = ulp_close(out["uv"][keep], ref["uv"][keep], 4) | (np.abs(out["uv"][keep] - ref["uv"][keep]) < 1e-15)
    assert ok.all(), f"{name}: uv differs on {(~ok).sum()} rays"


@pytest.mark.parametrize("mode", [abi.TRAVERSAL_FAST, abi.TRAVERSAL_MIRROR])
@pytest.mark.parametrize("name", ALL_SCENES)
def test_hits_vs_reference_golden(ctx, name, mode):
    """b200rt_trace on the golden ray batch == the reference's own <world>.hit records."""
    g = golden(name)
    ctx.set_option(abi.OPT_TRAVERSAL, mode)
    ctx.upload_scene(g.scene)
    out = ctx.trace(g["trace_rays"], seed=99)
    ctx.set_option(abi.OPT_TRAVERSAL, abi.TRAVERSAL_FAST)
    ref = {k: g["trace_" + k] for k in ("prim", "t", "uv", "front", "p", "normal")}
    keep = np.ones(len(out["prim"]), bool)
    if name in VOLUME_SCENES:
        vol = np.array(sorted(g.volume_prims()))
        keep = ~np.isin(out["prim"], vol) & ~np.isin(ref["prim"], vol)
    check_records(name, out, ref, keep)


@pytest.mark.parametrize("mode", [abi.TRAVERSAL_FAST, abi.TRAVERSAL_MIRROR])
@pytest.mark.parametrize("name", sorted(VOLUME_SCENES))
def test_volume_hits_vs_oracle(ctx, name, mode, build_oracle):
    """Same Philox free-flight draws on both sides: volume hits are comparable too (log() may differ in the last place)."""
    g = golden(name)
    ctx.set_option(abi.OPT_TRAVERSAL, mode)
    ctx.upload_scene(g.scene)
    out = ctx.trace(g["trace_rays"], seed=99)
    ctx.set_option(abi.OPT_TRAVERSAL, abi.TRAVERSAL_FAST)
    ref = build_oracle.trace(g.scene, g["trace_rays"], seed=99)
    assert (out["prim"] == ref["prim"]).mean() > 0.9999
    m = out["prim"] == ref["prim"]
    assert np.allclose(out["t"][m], ref["t"][m], rtol=1e-12, atol=0)
    assert np.isin(ref["prim"], np.array(sorted(g.volume_prims()))).sum() > 20


@pytest.mark.parametrize("name", ALL_SCENES)
def test_camera_rays_vs_oracle(ctx, name, build_oracle):
    g = golden(name)
    ctx.set_camera(g.cam_small)
    rays = ctx.generate_rays(31337, 1, 3)
    ref = build_oracle.generate_rays(g.cam_small, 31337, 1, 3)
    if g.cam_small.mode == abi.CAM_FISHEYE:
        assert (np.isnan(rays) == np.isnan(ref)).all()
        assert np.allclose(rays, ref, rtol=1e-13, atol=1e-15, equal_nan=True)
    else:
        assert same(rays, ref).all()


@pytest.mark.parametrize("name", ALL_SCENES)
def test_textures_vs_reference_golden(ctx, name):
    g = golden(name)
    ctx.upload_scene(g.scene)
    rows = g["tex_rows"]
    kinds = g.scene.textures["kind"]
    for t in np.unique(rows):
        rgb = ctx.texture_sample(int(t), g["tex_uvp"][rows == t])
        ref = g["tex_rgb"][rows == t]
        if kinds[t] in (abi.TEX_PERLIN, abi.TEX_WORLEY, abi.TEX_VORONOI):  # sin() inside: last-place differences allowed
            if kinds[t] == abi.TEX_PERLIN:
                assert np.allclose(rgb, ref, rtol=0, atol=1e-12), f"{name}: texture {t}"
            elif kinds[t] == abi.TEX_WORLEY:  # fract(sin(x) * 43758.5453) amplifies a 1-ulp difference of sin() by ~1e5, and a nearest-cell flip is possible
                assert (np.abs(rgb - ref) < 1e-8).mean() > 0.98, f"{name}: texture {t}"
            else:  # voronoi hashes the hashed position again (src/noise.h:193): a last-place sin() difference is amplified twice, ~1e-16 * 43758 * 312 * 43758
                assert (np.abs(rgb - ref) < 1e-3).mean() > 0.98, f"{name}: texture {t}"
                assert (np.abs(rgb - ref) < 1e-8).mean() > 0.3, f"{name}: texture {t}"
        else:
            assert same(rgb, ref).all(), f"{name}: texture {t}"


def outlier_bounds(name, a, b, fin, close, spp):
    """The pixels that do NOT agree to 1e-9 are those where one sample took another branch on a last-place libm difference (a
    scatter decision, a volume free-flight distance).  Such a pixel differs by at most one or two whole samples, i.e. by no more
    than 2 x (the brightest single sample of either image) / spp; the brightest sample is at most spp x the brightest pixel mean.
    Anything larger (a dropped subtree, a wrong tie-break that moves many samples of a pixel) fails here."""
    off = fin & ~close
    if not off.any():
        return
    brightest_sample = spp * max(np.abs(a[fin]).max(), np.abs(b[fin]).max())
    worst = np.abs(a - b)[off].max()
    print(f"[{name}] {int(off.sum())} of {int(fin.sum())} values differ beyond 1e-9; largest difference {worst:.3g} (brightest sample <= {brightest_sample:.3g})")
    assert worst <= 2.0 * brightest_sample / spp, f"{name}: an outlier differs by {worst}, more than two whole samples"
    # and they are isolated events, not a region: no pixel row holds more than a few of them
    rows = off.reshape(off.shape[0], -1).sum(axis=1)
    assert rows.max() <= max(6, 0.05 * off[0].size), f"{name}: {rows.max()} differing values in one image row"


def render_pair(ctx, oracle, g, seed, spp, mode=abi.TRAVERSAL_FAST, kill_zero=0):
    ctx.set_option(abi.OPT_TRAVERSAL, mode)
    ctx.set_option(abi.OPT_KILL_ZERO_WEIGHT, kill_zero)
    ctx.upload_scene(g.scene)
    ctx.set_camera(g.cam_small)
    img, st = ctx.render(seed, 0, spp)
    ctx.set_option(abi.OPT_TRAVERSAL, abi.TRAVERSAL_FAST)
    ctx.set_option(abi.OPT_KILL_ZERO_WEIGHT, 1)
    ref = oracle.render(g.scene, g.cam_small, seed, 0, spp, threads=0)
    return img, st, ref


@pytest.mark.parametrize("name", [s for s in ALL_SCENES if s not in ("motion_ball", "noise_voronoi")])  # voronoi: see test_textures_vs_reference_golden
def test_render_matches_oracle_sample_for_sample(ctx, name, build_oracle):
    """Same draws on both sides => the spp-sample image itself must agree, not just its expectation."""
    g = golden(name)
    spp = 8
    img, st, ref = render_pair(ctx, build_oracle, g, 4711, spp)
    a, b = img / spp, ref["sum"] / spp
    assert np.array_equal(np.isfinite(a), np.isfinite(b)), f"{name}: non-finite pixels differ from the oracle's"
    fin = np.isfinite(b)
    assert fin.mean() > 0.995
    close = np.abs(a - b) <= 1e-9 * np.maximum(np.abs(b), 1e-3)
    assert close[fin].mean() > 0.995, f"{name}: only {close[fin].mean():.4f} of pixels agree to 1e-9"
    outlier_bounds(name, a, b, fin, close, spp)
    rmse = np.sqrt(((a - b)[fin] ** 2).mean())
    assert rmse < 1e-3 * max(1.0, np.abs(b[fin]).max()), f"{name}: rmse {rmse}"
    assert abs(a[fin].mean() - b[fin].mean()) <= 0.005 * abs(b[fin].mean()) + 1e-12
    assert st["samples"] == ref["stats"]["samples"] and st["nan_samples"] == ref["stats"]["nan_samples"]
    assert abs(int(st["rays"]) - int(ref["stats"]["rays"])) <= 0.002 * ref["stats"]["rays"] + 2, (st["rays"], ref["stats"]["rays"])


@pytest.mark.parametrize("name", ["cornell", "rtnw", "fox", "cornell_volume"])
def test_render_mirror_mode_matches_oracle(ctx, name, build_oracle):
    g = golden(name)
    spp = 4
    img, st, ref = render_pair(ctx, build_oracle, g, 99, spp, mode=abi.TRAVERSAL_MIRROR)
    a, b = img / spp, ref["sum"] / spp
    assert np.array_equal(np.isfinite(a), np.isfinite(b)), f"{name}: non-finite pixels differ from the oracle's"
    fin = np.isfinite(b)
    close = np.abs(a - b) <= 1e-9 * np.maximum(np.abs(b), 1e-3)
    assert close[fin].mean() > 0.995
    outlier_bounds(name, a, b, fin, close, spp)


@pytest.mark.parametrize("name", CONFIG_SCENES)
def test_converged_image_vs_reference(ctx, name):
    """North-star image criterion against the reference's OWN render (golden sums): per-pixel z-scores consistent with
    Monte-Carlo error, image mean within 0.5 % (+ standard error)."""
    from conftest import z_scores

    g = golden(name)
    n_ref = int(g["ref_spp"])
    ctx.upload_scene(g.scene)
    ctx.set_camera(g.cam_small)
    H, W = g.cam_small.image_height, g.cam_small.image_width
    n = 4 * n_ref
    m = n // 16
    s = np.zeros((H, W, 3))
    batch_means = []
    for k in range(0, n, m):  # 16 batches: the spread of their means gives the per-sample variance (the ABI returns sums only)
        part, _ = ctx.render(5, k, k + m)
        s += part
        batch_means.append(part / m)
    fin = np.isfinite(s) & np.isfinite(g["ref_sum"])
    mean_g, mean_r = s / n, g["ref_sum"] / n_ref
    var_r = np.maximum(g["ref_sumsq"] / n_ref - mean_r**2, 0)
    with np.errstate(invalid="ignore"):
        var_g = np.array(batch_means).var(axis=0, ddof=1) * m
    var = np.where(fin, (n * var_g + n_ref * var_r) / (n + n_ref), 0)  # pooled, as conftest.z_scores
    se = np.sqrt(var / n + var / n_ref)
    se = np.maximum(se, 1e-9 * np.maximum(np.abs(mean_r), 1e-3))  # constant pixels (solid background): rounding only
    with np.errstate(divide="ignore", invalid="ignore"):
        z = np.where(fin, (mean_g - mean_r) / se, 0)
    assert np.sqrt((z**2).mean()) < 1.3 and np.abs(z).max() < 7, (np.sqrt((z**2).mean()), np.abs(z).max())
    tot_g, tot_r = mean_g[fin].mean(), mean_r[fin].mean()
    se_tot = np.sqrt((var / n + var / n_ref)[fin].sum()) / fin.sum()
    assert abs(tot_g - tot_r) <= 0.005 * abs(tot_r) + 5 * se_tot


def test_sample_ranges_add_up_and_wave_size_is_irrelevant(ctx):
    """Properties the multi-GPU split relies on: render[0,b) == render[0,a) + render[a,b) and the result does not depend
    on the wave size (only fp64 atomic summation order differs)."""
    g = golden("rtow")
    ctx.upload_scene(g.scene)
    ctx.set_camera(g.cam_small)
    whole, st = ctx.render(1234, 0, 24)
    a, _ = ctx.render(1234, 0, 10)
    b, _ = ctx.render(1234, 10, 24)
    assert np.allclose(whole, a + b, rtol=1e-12, atol=1e-12)
    ctx.set_option(abi.OPT_WAVE_PATHS, 4096)
    small, st2 = ctx.render(1234, 0, 24)
    ctx.set_option(abi.OPT_WAVE_PATHS, 1 << 25)
    assert np.allclose(whole, small, rtol=1e-12, atol=1e-12)
    assert st["rays"] == st2["rays"] and st["samples"] == st2["samples"] == 24 * g.cam_small.image_width * g.cam_small.image_height


@pytest.mark.parametrize("name", ["cornell", "rtow", "rtnw", "fox", "cornell_volume", "cornell_specular", "three_ball_lens", "light_earth"])
def test_tail_kernel_does_not_change_the_image(ctx, name):
    """B200RT_OPT_TAIL_PATHS: finishing the last paths of a frame in one kernel (k_tail) instead of one wave per bounce uses the
    same functions and the same draws: sums equal up to fp64 atomic order, rays and samples counted exactly the same.  2^20
    sends the whole small frame through k_tail as soon as its samples are generated; 0 switches it off."""
    g = golden(name)
    ctx.upload_scene(g.scene)
    ctx.set_camera(g.cam_small)
    try:
        ctx.set_option(abi.OPT_TAIL_PATHS, 0)
        base, st = ctx.render(2024, 0, 12)
        for tail in (1 << 20, 32768, 600):
            ctx.set_option(abi.OPT_TAIL_PATHS, tail)
            img, st2 = ctx.render(2024, 0, 12)
            assert (st2["samples"], st2["nan_samples"], st2["rays"]) == (st["samples"], st["nan_samples"], st["rays"]), tail
            assert np.allclose(np.nan_to_num(base), np.nan_to_num(img), rtol=1e-12, atol=1e-12), tail
            if tail == 1 << 20:
                assert st2["waves"] < st["waves"]
    finally:
        ctx.set_option(abi.OPT_TAIL_PATHS, 32768)


@pytest.mark.parametrize("name", ["fox", "skybox_fisheye"])
def test_fisheye_samples_outside_the_image_circle_are_dropped_not_lost(ctx, name):
    """Fisheye cameras: k_generate<COMPACT> drops the samples that produce no ray (src/camera.h:268-282, NaN direction) and tops
    the wave up in a second pass.  Small waves force many partially filled waves: every sample must still be counted exactly
    once (samples, nan_samples, rays) and the sums must not depend on the wave size or on how the range is cut."""
    g = golden(name)
    ctx.upload_scene(g.scene)
    ctx.set_camera(g.cam_small)
    npix = g.cam_small.image_width * g.cam_small.image_height
    whole, st = ctx.render(99, 0, 20)
    assert st["samples"] == 20 * npix and 0 < st["nan_samples"] < st["samples"]
    try:
        for wave in (1024, 5000, 1 << 16):
            ctx.set_option(abi.OPT_WAVE_PATHS, wave)
            img, st2 = ctx.render(99, 0, 20)
            assert (st2["samples"], st2["nan_samples"], st2["rays"]) == (st["samples"], st["nan_samples"], st["rays"]), wave
            assert np.allclose(np.nan_to_num(whole), np.nan_to_num(img), rtol=1e-12, atol=1e-12), wave
        a, sa = ctx.render(99, 0, 7)
        b, sb = ctx.render(99, 7, 20)
        assert sa["nan_samples"] + sb["nan_samples"] == st["nan_samples"] and sa["rays"] + sb["rays"] == st["rays"]
        assert np.allclose(np.nan_to_num(whole), np.nan_to_num(a + b), rtol=1e-12, atol=1e-12)
    finally:
        ctx.set_option(abi.OPT_WAVE_PATHS, 1 << 25)


@pytest.mark.parametrize("name", ["rtow", "rtnw", "fox", "cornell_volume", "three_ball_lens", "skybox_fisheye"])
def test_traversal_tuning_options_do_not_change_the_image(ctx, name):
    """Lane refill and speculative (postponed-leaf) traversal change the order in which leaves are tested, never the hit:
    same rays, same samples, same sums (up to fp64 atomic order)."""
    g = golden(name)
    ctx.upload_scene(g.scene)
    ctx.set_camera(g.cam_small)
    base, st = ctx.render(77, 0, 6)
    try:
        for opts in ({abi.OPT_SPECULATIVE_LEAF: 1}, {abi.OPT_REFILL_LANES: 12}, {abi.OPT_SPECULATIVE_LEAF: 1, abi.OPT_REFILL_LANES: 20}, {abi.OPT_FUSE_MISS: 0},
                     {abi.OPT_EXTEND_BLOCKS: 9}, {abi.OPT_EXTEND_STAGES: 1}, {abi.OPT_SMEM_CARVEOUT: 100}, {abi.OPT_WAVE_LOOP: 0},
                     {abi.OPT_EXTEND_LOOP: 1}, {abi.OPT_EXTEND_LOOP: 1, abi.OPT_LEAF_LANES: 1}, {abi.OPT_EXTEND_LOOP: 1, abi.OPT_LEAF_LANES: 32, abi.OPT_REFILL_LANES: 8},
                     {abi.OPT_EXTEND_LOOP: 1, abi.OPT_BVH_WIDTH: 4}, {abi.OPT_L2_PERSIST: 0}, {abi.OPT_TRI_F32: 0}):
            for k, v in opts.items():
                ctx.set_option(k, v)
            if abi.OPT_BVH_WIDTH in opts or abi.OPT_TRI_F32 in opts:  # take effect at upload
                ctx.upload_scene(g.scene)
            img, st2 = ctx.render(77, 0, 6)
            ctx.set_option(abi.OPT_SPECULATIVE_LEAF, 0)
            ctx.set_option(abi.OPT_REFILL_LANES, 0)
            ctx.set_option(abi.OPT_FUSE_MISS, 1)
            ctx.set_option(abi.OPT_EXTEND_BLOCKS, 4)
            ctx.set_option(abi.OPT_EXTEND_STAGES, 2)
            ctx.set_option(abi.OPT_SMEM_CARVEOUT, -1)
            ctx.set_option(abi.OPT_WAVE_LOOP, 1)
            ctx.set_option(abi.OPT_EXTEND_LOOP, -1)
            ctx.set_option(abi.OPT_LEAF_LANES, 0)
            ctx.set_option(abi.OPT_TRI_F32, 1)
            ctx.set_option(abi.OPT_L2_PERSIST, 1)
            if abi.OPT_BVH_WIDTH in opts or abi.OPT_TRI_F32 in opts:
                ctx.set_option(abi.OPT_BVH_WIDTH, 0)
                ctx.upload_scene(g.scene)
            assert st2["rays"] == st["rays"], opts
            fin = np.isfinite(base)
            assert np.array_equal(fin, np.isfinite(img))
            assert np.allclose(base[fin], img[fin], rtol=1e-12, atol=1e-12), opts
    finally:
        ctx.set_option(abi.OPT_SPECULATIVE_LEAF, 0)
        ctx.set_option(abi.OPT_REFILL_LANES, 0)
        ctx.set_option(abi.OPT_FUSE_MISS, 1)
        ctx.set_option(abi.OPT_EXTEND_BLOCKS, 4)
        ctx.set_option(abi.OPT_EXTEND_STAGES, 2)
        ctx.set_option(abi.OPT_SMEM_CARVEOUT, -1)
        ctx.set_option(abi.OPT_WAVE_LOOP, 1)
        ctx.set_option(abi.OPT_EXTEND_LOOP, -1)
        ctx.set_option(abi.OPT_LEAF_LANES, 0)
        ctx.set_option(abi.OPT_BVH_WIDTH, 0)


def test_white_furnace_known_answer(ctx):
    """White Sphere (src/main.cc:500-510): albedo-1 metal under a white background: every sample is 1 unless the path
    runs out of depth re-hitting the sphere."""
    g = golden("white_sphere")
    ctx.upload_scene(g.scene)
    ctx.set_camera(g.cam_small)
    img, st = ctx.render(3, 0, 64)
    m = img / 64
    assert m.max() <= 1.0 + 1e-12 and m.mean() > 0.99


def test_finalize_matches_write_color(ctx):
    g = golden("cornell")
    ctx.set_camera(g.cam_small)
    H, W = g.cam_small.image_height, g.cam_small.image_width
    rng = np.random.RandomState(0)
    s = rng.rand(H, W, 3) * 40
    s[0, 0, 0] = np.nan
    s[0, 1, 1] = -3.0
    s[0, 2, 2] = 1e300
    lin, ppm = ctx.finalize(s, 16, want_ppm=True)
    assert same(lin, s / 16).all()
    c = s / 16
    with np.errstate(invalid="ignore"):
        expect = np.where(c > 0, np.power(np.where(c > 0, c, 1.0), 1 / 2.2), 0.0) * 255.999
    big = ~(np.abs(expect) < 2**31)
    ref = np.where(big, -(2**31), np.where(big, 0, expect).astype(np.int64))
    assert (np.abs(ppm.astype(np.int64) - ref) <= 1).all() and (ppm[big] == -(2**31)).all()


def test_format_ppm_is_the_reference_file_image(ctx):
    """b200rt_format_ppm == header + one `out << r << ' ' << g << ' ' << b << '\\n'` per pixel (src/camera.h:149-151, src/color.h:35),
    with the integers of b200rt_finalize: negative, > 255 and INT_MIN values print like operator<<(int)."""
    g = golden("cornell")
    for W, H in ((37, 29), (64, 33), (1, 1), (1030, 3)):
        cam = g.cam_small.copy()
        cam.image_width, cam.image_height = W, H
        ctx.set_camera(cam)
        rng = np.random.RandomState(W)
        s = rng.rand(H, W, 3) * 10 ** rng.uniform(-3, 3, size=(H, W, 3))
        s.reshape(-1)[:3] = [np.nan, -3.0, 1e300]
        s.reshape(-1)[-1] = 0.0
        _, ppm = ctx.finalize(s, 16, want_ppm=True)
        want = f"P3\n{W} {H}\n255\n" + "".join(f"{r} {gg} {b}\n" for r, gg, b in ppm.reshape(-1, 3).tolist())
        got = ctx.format_ppm(s, 16)
        assert got == want.encode()
        with pytest.raises(pkg.backend.B200RTError):
            ctx.format_ppm(s, 16, capacity=len(want) - 1)
        assert ctx.format_ppm(s, 16, capacity=len(want)) == want.encode()
    ctx.set_camera(g.cam_small)


def test_errors_are_loud(ctx):
    from conftest import Scene

    c2 = pkg.backend.Context(0)
    with pytest.raises(pkg.backend.B200RTError):
        c2.trace(np.zeros((4, 7)))  # no scene yet
    bad = Scene.load(golden("cornell").z)
    bad.quad_mat = bad.quad_mat.copy()
    bad.quad_mat[0] = 999
    with pytest.raises(pkg.backend.B200RTError):
        c2.upload_scene(bad)
    with pytest.raises(pkg.backend.B200RTError):
        pkg.backend.Context(99)
    c2.close()


def test_sponza_standin_hits(ctx, build_oracle):
    """BASELINE.json configs[4] (stand-in mesh, 262 267 triangles; the real Sponza.bin is absent from the reference checkout):
    compacted traversal == node-for-node mirror of the x-sorted reference tree on the GPU for a full batch of camera rays,
    and both == the CPU oracle on a subset (the reference tree costs ~1 ms per ray on a CPU core)."""
    sc = pkg.scene.sponza_standin_scene()
    cam = pkg.camera.initialize_perspective(192, 16.0 / 9.0, (500, 320, 90), (0, 280, 0), 1, 45.0, 4, 5)
    ctx.upload_scene(sc)
    ctx.set_camera(cam)
    rays = ctx.generate_rays(9, 0, 2)
    hit = ctx.trace(rays)
    sec = rays[hit["prim"] >= 0].copy()  # secondary rays from the first hits
    rng = np.random.RandomState(2)
    sec[:, 0:3] = hit["p"][hit["prim"] >= 0]
    sec[:, 3:6] = rng.normal(size=(len(sec), 3))
    batch = np.concatenate([rays, sec])
    fast = ctx.trace(batch)
    ctx.set_option(abi.OPT_TRAVERSAL, abi.TRAVERSAL_MIRROR)
    ctx.upload_scene(sc)
    mir = ctx.trace(batch)
    ctx.set_option(abi.OPT_TRAVERSAL, abi.TRAVERSAL_FAST)
    for k in ("prim", "front", "t", "p", "normal"):
        assert same(fast[k], mir[k]).all(), k
    assert (fast["prim"] >= 0).mean() > 0.7
    sub = batch[:: max(1, len(batch) // 1500)]
    ref = build_oracle.trace(sc, sub)
    got = ctx.trace(sub)
    for k in ("prim", "front", "t", "p", "normal"):
        assert same(got[k], ref[k]).all(), k
    img, st = ctx.render(3, 0, 2)
    assert np.isfinite(img).all() and st["rays"] > st["samples"]


@pytest.mark.parametrize("builder", [abi.BVH_DEVICE_LBVH, abi.BVH_DEVICE_PLOC])
@pytest.mark.parametrize("name", ALL_SCENES)
def test_device_built_bvh_gives_the_same_hits(name, builder):
    """B200RT_OPT_BVH_BUILDER = DEVICE_LBVH (csrc/bvh_gpu.cuh) / DEVICE_PLOC (csrc/bvh_ploc.cuh + finish_tree): a different tree,
    the same records as the reference."""
    g = golden(name)
    c = pkg.backend.Context(0)
    try:
        c.set_option(abi.OPT_BVH_BUILDER, builder)
        if builder == abi.BVH_DEVICE_PLOC:
            c.set_option(abi.OPT_BVH_LEAF_SIZE, 2)  # small scenes default to one primitive per leaf: exercise the leaf-pair pass
        c.upload_scene(g.scene)
        info = c.bvh_info()
        if builder == abi.BVH_DEVICE_LBVH:
            assert info["n_leaves"] == info["n_prims"] and info["n_nodes"] == max(0, info["n_prims"] - 1)
        else:
            assert info["n_prims"] / 2 <= info["n_leaves"] <= info["n_prims"]
        assert 1 <= info["max_depth"] <= 64
        out = c.trace(g["trace_rays"], seed=99)
        ref = {k: g["trace_" + k] for k in ("prim", "t", "uv", "front", "p", "normal")}
        keep = np.ones(len(out["prim"]), bool)
        if name in VOLUME_SCENES:
            vol = np.array(sorted(g.volume_prims()))
            keep = ~np.isin(out["prim"], vol) & ~np.isin(ref["prim"], vol)
        check_records(name, out, ref, keep)
        # and the wavefront renders the same image from it (same samples, same hits; fp64 atomic order aside)
        c.set_camera(g.cam_small)
        img, st = c.render(5, 0, 4)
        c.set_option(abi.OPT_BVH_BUILDER, abi.BVH_HOST_SAH)
        c.upload_scene(g.scene)
        img0, st0 = c.render(5, 0, 4)
        assert st["rays"] == st0["rays"]
        fin = np.isfinite(img0)
        assert np.array_equal(fin, np.isfinite(img)) and np.allclose(img[fin], img0[fin], rtol=1e-12, atol=1e-12)
    finally:
        c.close()


def test_device_built_bvh_sponza_standin():
    sc = pkg.scene.sponza_standin_scene()
    cam = pkg.camera.initialize_perspective(160, 16.0 / 9.0, (500, 320, 90), (0, 280, 0), 1, 45.0, 4, 5)
    c = pkg.backend.Context(0)
    try:
        c.upload_scene(sc)
        c.set_camera(cam)
        rays = c.generate_rays(9, 0, 2)
        host = c.trace(rays)
        hinfo = c.bvh_info()
        c.set_option(abi.OPT_BVH_BUILDER, abi.BVH_DEVICE_LBVH)
        c.upload_scene(sc)
        dev = c.trace(rays)
        dinfo = c.bvh_info()
        for k in ("prim", "front", "t", "p", "normal"):
            assert same(host[k], dev[k]).all(), k
        assert dinfo["n_prims"] == hinfo["n_prims"] and dinfo["build_ms"] < hinfo["build_ms"]
        print(f"sponza stand-in BVH: host SAH {hinfo['build_ms']:.1f} ms (cost {hinfo['sah_cost']:.1f}), device LBVH {dinfo['build_ms']:.2f} ms (cost {dinfo['sah_cost']:.1f}, depth {dinfo['max_depth']})")
        c.set_option(abi.OPT_BVH_BUILDER, abi.BVH_DEVICE_PLOC)
        c.upload_scene(sc)
        ploc = c.trace(rays)
        pinfo = c.bvh_info()
        for k in ("prim", "front", "t", "p", "normal"):
            assert same(host[k], ploc[k]).all(), k
        assert pinfo["n_prims"] == hinfo["n_prims"] and pinfo["build_ms"] < hinfo["build_ms"]
        assert pinfo["n_leaves"] < pinfo["n_prims"]  # leaf pairs were formed
        c.upload_scene(sc)  # node indices come from a scan, not from atomics: the same scene gives the same tree
        again = c.bvh_info()
        for k in ("n_nodes", "n_leaves", "max_depth", "sah_cost"):
            assert again[k] == pinfo[k], k
        print(f"                     device PLOC + host finish {pinfo['build_ms']:.2f} ms (cost {pinfo['sah_cost']:.1f}, depth {pinfo['max_depth']}, {pinfo['n_nodes']} BVH4 nodes vs {hinfo['n_nodes']})")
    finally:
        c.close()


@pytest.mark.parametrize("name", ["cornell", "rtow", "three_ball_lens"])
def test_reference_side_binding_end_to_end(name, tmp_path):
    """INTEGRATION.md's drop-in: b200::render(cam, out, world, light) of include/b200rt_adaptor.hpp, compiled against the
    UNMODIFIED reference headers (oracle/_ref), called on live reference objects in place of cam.render(...): the P3 file
    it writes and the cam.image_ it leaves equal what the C ABI gives for the flattened scene on the same Philox seed."""
    import refharness

    if not refharness.available():
        pytest.skip("oracle/_ref/libref_harness.so not built (needs /root/reference at build time)")
    rs = refharness.RefScene(name, width=56, spp=5, depth=12)
    path = str(tmp_path / "b200.ppm")
    secs, img = rs.render_b200(seed=0x5EED, ppm_path=path)
    text = open(path, "rb").read()
    H, W = rs.camera.image_height, rs.camera.image_width
    assert text.startswith(f"P3\n{W} {H}\n255\n".encode())
    with pkg.backend.Context(0) as c:
        c.upload_scene(rs.scene)
        c.set_camera(rs.camera)
        s, _ = c.render(0x5EED, 0, rs.camera.samples_per_pixel)
        lin = c.finalize(s, rs.camera.samples_per_pixel)
        want = c.format_ppm(s, rs.camera.samples_per_pixel)
    fin = np.isfinite(lin)
    assert np.allclose(img[fin], lin[fin], rtol=1e-12, atol=1e-12)
    got_ints = np.array(text.split()[4:], dtype=np.int64)
    want_ints = np.array(want.split()[4:], dtype=np.int64)
    assert got_ints.shape == want_ints.shape == (H * W * 3,)
    assert (np.abs(got_ints - want_ints) <= 1).all() and (got_ints != want_ints).mean() < 1e-3  # fp64 atomic order can move a value across an integer boundary
    rs.close()


@pytest.mark.parametrize("name", ALL_SCENES)
def test_bvh4_gives_the_same_hits(name):
    """B200RT_OPT_BVH_WIDTH = 4 (the BVH2 collapsed into 128-byte four-child nodes): a different walk, the same records."""
    g = golden(name)
    c = pkg.backend.Context(0)
    try:
        c.set_option(abi.OPT_BVH_WIDTH, 4)
        c.upload_scene(g.scene)
        out = c.trace(g["trace_rays"], seed=99)
        ref = {k: g["trace_" + k] for k in ("prim", "t", "uv", "front", "p", "normal")}
        keep = np.ones(len(out["prim"]), bool)
        if name in VOLUME_SCENES:
            vol = np.array(sorted(g.volume_prims()))
            keep = ~np.isin(out["prim"], vol) & ~np.isin(ref["prim"], vol)
        check_records(name, out, ref, keep)
        c.set_camera(g.cam_small)
        img, st = c.render(5, 0, 4)
        c.set_option(abi.OPT_BVH_WIDTH, 2)
        c.upload_scene(g.scene)
        img0, st0 = c.render(5, 0, 4)
        assert st["rays"] == st0["rays"]
        fin = np.isfinite(img0)
        assert np.array_equal(fin, np.isfinite(img)) and np.allclose(img[fin], img0[fin], rtol=1e-12, atol=1e-12)
    finally:
        c.close()


def _empty_world(g):
    """A hittable_list with no objects (src/hittable_list.h:20-31 returns false for every ray) under the scene's background."""
    import copy

    s = copy.copy(g.scene)
    s.nodes = g.scene.nodes[:1].copy()
    s.nodes["kind"][0], s.nodes["a"][0], s.nodes["b"][0] = abi.NODE_LIST, 0, 0
    s.list_children = g.scene.list_children[:0].copy()
    s.root = 0
    s.n_primitives = 0
    return s


@pytest.mark.parametrize("builder", [abi.BVH_HOST_SAH, abi.BVH_DEVICE_LBVH, abi.BVH_DEVICE_PLOC])
@pytest.mark.parametrize("width", [2, 4])
def test_edge_cases_empty_and_single_primitive_worlds(builder, width, build_oracle):
    """Empty world: every ray misses, every sample is the background (here solid white: sum == spp).  One-primitive world:
    the root of the compacted tree is a leaf.  Empty ray batches and empty sample ranges are no-ops.  All against the oracle."""
    g = golden("white_sphere")
    with pkg.backend.Context(0) as c:
        c.set_option(abi.OPT_BVH_BUILDER, builder)
        c.set_option(abi.OPT_BVH_WIDTH, width)
        # one primitive
        c.upload_scene(g.scene)
        info = c.bvh_info()
        assert info["n_prims"] == 1 and info["n_leaves"] == 1
        out, ref = c.trace(g["trace_rays"], seed=99), build_oracle.trace(g.scene, g["trace_rays"], seed=99)
        for k in ("prim", "t", "p", "normal", "front"):
            assert same(out[k], ref[k]).all(), k
        assert (out["prim"] >= 0).any() and (out["prim"] < 0).any()
        c.set_camera(g.cam_small)
        img, st = c.render(3, 0, 4)
        want = build_oracle.render(g.scene, g.cam_small, 3, 0, 4)
        assert np.allclose(img, want["sum"], rtol=1e-9, atol=1e-12) and st["rays"] == want["stats"]["rays"]
        # nothing to do
        assert len(c.trace(np.zeros((0, 7)))["prim"]) == 0
        z, stz = c.render(3, 5, 5)
        assert not z.any() and stz["samples"] == 0 and stz["rays"] == 0
        # empty world
        e = _empty_world(g)
        c.upload_scene(e)
        assert c.bvh_info()["n_prims"] == 0
        out = c.trace(g["trace_rays"], seed=99)
        assert (out["prim"] == -1).all()
        img, st = c.render(3, 0, 4)
        want = build_oracle.render(e, g.cam_small, 3, 0, 4)
        assert np.allclose(img, want["sum"], rtol=1e-12, atol=0) and st["rays"] == st["samples"] == want["stats"]["rays"]


@pytest.mark.parametrize("name", ["rtow", "cornell", "rtnw"])
def test_two_pipelines_render_the_same_frame(name):
    """B200RT_OPT_PIPELINES = 2: two independent wavefronts over the two halves of the sample range, on two streams, adding
    into the same accumulator: same samples, same rays, same sums (fp64 atomic order aside)."""
    g = golden(name)
    cam = g.cam_full.copy()
    cam.samples_per_pixel = 8
    with pkg.backend.Context(0) as c:
        c.upload_scene(g.scene)
        c.set_camera(cam)
        c.set_option(abi.OPT_PIPELINES, 1)
        one, st1 = c.render(9, 3, 11)
        for ctas, npipe in ((4, 2), (3, 2), (2, 2), (4, 3), (4, 4)):
            c.set_option(abi.OPT_PIPELINES, npipe)
            c.set_option(abi.OPT_PIPE_EXTEND_CTAS, ctas)
            two, st2 = c.render(9, 3, 11)
            assert st2["rays"] == st1["rays"] and st2["samples"] == st1["samples"]
            fin = np.isfinite(one)
            assert np.array_equal(fin, np.isfinite(two)) and np.allclose(one[fin], two[fin], rtol=1e-11, atol=1e-11)
        odd, _ = c.render(9, 0, 1)  # a one-sample range cannot be split: falls back to one pipeline
        c.set_option(abi.OPT_PIPELINES, 1)
        ref, _ = c.render(9, 0, 1)
        c.set_option(abi.OPT_PIPELINES, 2)
        assert np.allclose(odd[np.isfinite(ref)], ref[np.isfinite(ref)], rtol=1e-12, atol=1e-12)
