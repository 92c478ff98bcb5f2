"""ctypes front end of oracle/_ref/libref_harness.so — TEST INFRASTRUCTURE.

The shared object is the UNMODIFIED reference (compiled from /root/reference/src by oracle/Makefile)
plus oracle/ref_harness.cc.  Only tests/, __graft_entry__.smoke() and the cpu_baseline /
--impl reference legs of bench.py may import this module.
"""
import ctypes as C
import importlib
import os
import sys

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_ROOT = os.path.dirname(_HERE)
if _ROOT not in sys.path:
    sys.path.insert(0, _ROOT)
_pkg = importlib.import_module("cpu-ray-tracing-implementation_b200")
abi = _pkg._abi
Scene = _pkg.scene.Scene

LIB_PATH = os.path.join(_HERE, "_ref", "libref_harness.so")
REFERENCE_ASSETS = "/root/reference/assets"


class RenderStats(C.Structure):
    _fields_ = [
        ("samples", C.c_uint64),
        ("rays", C.c_uint64),
        ("nan_samples", C.c_uint64),
        ("seconds", C.c_double),
        ("threads", C.c_int),
        ("rng_mode", C.c_int),
    ]


def available():
    return os.path.exists(LIB_PATH)


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not available():
            raise RuntimeError(f"{LIB_PATH} missing: run `make -f oracle/Makefile ref` where /root/reference exists")
        # the harness also compiles the reference-side binding (b200::render of include/b200rt_adaptor.hpp); the product
        # library is loaded by ref_render_b200 itself, only when that entry is called (never by the CPU reference arm)
        L = C.CDLL(LIB_PATH)
        vp, u32, i32, sz, dbl = C.c_void_p, C.c_uint, C.c_int, C.c_size_t, C.c_double
        L.ref_last_error.restype = C.c_char_p
        L.ref_scene_create.argtypes = [C.c_char_p, i32, i32, i32, C.c_char_p, u32, vp, i32]
        L.ref_scene_create.restype = i32
        L.ref_scene_destroy.argtypes = [i32]
        L.ref_scene_desc.argtypes = [i32]
        L.ref_scene_desc.restype = vp
        L.ref_scene_camera.argtypes = [i32, C.POINTER(abi.Camera)]
        L.ref_generate_rays.argtypes = [i32, u32, i32, vp]
        L.ref_trace.argtypes = [i32, vp, sz, dbl, dbl, u32, vp, vp, vp, vp, vp, vp]
        L.ref_texture_sample.argtypes = [i32, i32, vp, sz, vp]
        L.ref_ray_color.argtypes = [i32, vp, sz, i32, u32, vp]
        L.ref_render.argtypes = [i32, u32, i32, i32, i32, vp, vp, C.POINTER(RenderStats)]
        L.ref_render_stock.argtypes = [i32, u32, i32, C.c_char_p, vp]
        L.ref_render_stock.restype = dbl
        L.ref_pstl_is_parallel.restype = i32
        L.ref_render_b200.argtypes = [i32, C.c_uint64, i32, i32, C.c_char_p, vp, C.c_char_p]
        L.ref_render_b200.restype = dbl
        _lib = L
    return _lib


class RefScene:
    """A scene built through the reference's own constructors."""

    def __init__(self, name, width=0, spp=0, depth=0, assets=REFERENCE_ASSETS, construct_seed=1, mesh=None):
        L = lib()
        mesh_p, n_mesh = None, 0
        if mesh is not None:
            self._mesh = np.ascontiguousarray(mesh, dtype=np.float32).reshape(-1, 9)
            mesh_p, n_mesh = self._mesh.ctypes.data, len(self._mesh)
        self.h = L.ref_scene_create(name.encode(), width, spp, depth, assets.encode(), construct_seed, mesh_p, n_mesh)
        if self.h < 0:
            raise RuntimeError(f"ref_scene_create({name}): {L.ref_last_error().decode()}")
        self.name = name
        self.scene = Scene.from_desc_pointer(L.ref_scene_desc(self.h))
        self.scene.name = name
        self.camera = abi.Camera()
        L.ref_scene_camera(self.h, C.byref(self.camera))

    def close(self):
        if self.h > 0:
            lib().ref_scene_destroy(self.h)
            self.h = -1

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def npix(self):
        return self.camera.image_width * self.camera.image_height

    def generate_rays(self, seed, spp=1):
        rays = np.zeros((spp * self.npix, 7), dtype=np.float64)
        assert lib().ref_generate_rays(self.h, seed, spp, rays.ctypes.data) == 0
        return rays

    def trace(self, rays, tmin=0.001, tmax=np.inf, seed=1):
        rays = np.ascontiguousarray(rays, dtype=np.float64).reshape(-1, 7)
        n = len(rays)
        out = dict(
            prim=np.zeros(n, np.int32),
            t=np.zeros(n, np.float64),
            uv=np.zeros((n, 2), np.float64),
            front=np.zeros(n, np.uint8),
            p=np.zeros((n, 3), np.float64),
            normal=np.zeros((n, 3), np.float64),
        )
        rc = lib().ref_trace(self.h, rays.ctypes.data, n, tmin, tmax, seed, *(out[k].ctypes.data for k in ("prim", "t", "uv", "front", "p", "normal")))
        assert rc == 0
        return out

    def texture_sample(self, tex, uvp):
        uvp = np.ascontiguousarray(uvp, dtype=np.float64).reshape(-1, 5)
        rgb = np.zeros((len(uvp), 3), np.float64)
        assert lib().ref_texture_sample(self.h, tex, uvp.ctypes.data, len(uvp), rgb.ctypes.data) == 0
        return rgb

    def ray_color(self, rays, depth, seed=1):
        rays = np.ascontiguousarray(rays, dtype=np.float64).reshape(-1, 7)
        rgb = np.zeros((len(rays), 3), np.float64)
        assert lib().ref_ray_color(self.h, rays.ctypes.data, len(rays), depth, seed, rgb.ctypes.data) == 0
        return rgb

    def render(self, spp, seed=12345, threads=0, rng_mode=1, want_sums=True):
        """-> (sum, sumsq, stats dict).  rng_mode 1 = thread-local streams on `threads` cores (0 = all)."""
        if threads <= 0:
            threads = os.cpu_count() or 1
        n = self.npix * 3
        s = np.zeros(n, np.float64) if want_sums else None
        q = np.zeros(n, np.float64) if want_sums else None
        st = RenderStats()
        rc = lib().ref_render(self.h, seed, spp, threads, rng_mode, s.ctypes.data if want_sums else None, q.ctypes.data if want_sums else None, C.byref(st))
        assert rc == 0
        stats = {k: getattr(st, k) for k, _ in RenderStats._fields_}
        H, W = self.camera.image_height, self.camera.image_width
        if want_sums:
            return s.reshape(H, W, 3), q.reshape(H, W, 3), stats
        return None, None, stats

    def render_stock(self, spp=0, seed=1, ppm_path=None):
        """camera::render as written. -> (seconds, linear image)"""
        img = np.zeros(self.npix * 3, np.float64)
        secs = lib().ref_render_stock(self.h, seed, spp, ppm_path.encode() if ppm_path else None, img.ctypes.data)
        return secs, img.reshape(self.camera.image_height, self.camera.image_width, 3)


def _render_b200(self, spp=0, seed=0x5EED, device=0, ppm_path=None):
    """b200::render(cam, out, world, light): the drop-in binding in place of camera::render. -> (seconds, linear image)"""
    img = np.zeros(self.npix * 3, np.float64)
    product = os.path.join(_ROOT, "cpu-ray-tracing-implementation_b200", "libb200rt.so")
    secs = lib().ref_render_b200(self.h, seed, spp, device, ppm_path.encode() if ppm_path else None, img.ctypes.data, product.encode())
    if secs < 0:
        raise RuntimeError("ref_render_b200: " + lib().ref_last_error().decode())
    return secs, img.reshape(self.camera.image_height, self.camera.image_width, 3)


RefScene.render_b200 = _render_b200


def _generate_rays_logged(self, seed, spp=1, max_draws=40):
    n = spp * self.npix
    rays = np.zeros((n, 7), np.float64)
    draws = np.zeros((n, max_draws), np.float64)
    counts = np.zeros(n, np.int32)
    L = lib()
    L.ref_generate_rays_logged.argtypes = [C.c_int, C.c_uint, C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
    assert L.ref_generate_rays_logged(self.h, seed, spp, rays.ctypes.data, draws.ctypes.data, max_draws, counts.ctypes.data) == 0
    return rays, draws, counts


RefScene.generate_rays_logged = _generate_rays_logged


def _scatter_logged(self, rays, seed=5, max_draws=8):
    """One level of ray_color below the hit test through the reference's own objects (ref_scatter_logged).
    -> dict(rec12, out16, draws, counts); out16[:, 0] == -1 where the ray missed."""
    rays = np.ascontiguousarray(rays, dtype=np.float64).reshape(-1, 7)
    n = len(rays)
    rec = np.zeros((n, 12), np.float64)
    out = np.zeros((n, 16), np.float64)
    draws = np.zeros((n, max_draws), np.float64)
    counts = np.zeros(n, np.int32)
    L = lib()
    L.ref_scatter_logged.argtypes = [C.c_int, C.c_void_p, C.c_size_t, C.c_uint, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
    rc = L.ref_scatter_logged(self.h, rays.ctypes.data, n, seed, rec.ctypes.data, out.ctypes.data, draws.ctypes.data, max_draws, counts.ctypes.data)
    if rc != 0:
        raise RuntimeError("ref_scatter_logged: " + L.ref_last_error().decode())
    assert counts.max(initial=0) <= max_draws
    return dict(rec12=rec, out16=out, draws=draws, counts=counts)


RefScene.scatter_logged = _scatter_logged


def write_colors(rgb):
    """P3 text of an (H, W, 3) linear image through the reference's own write_color (src/color.h:22-36) -> bytes."""
    rgb = np.ascontiguousarray(rgb, dtype=np.float64)
    H, W = rgb.shape[:2]
    L = lib()
    L.ref_write_colors.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_size_t]
    L.ref_write_colors.restype = C.c_size_t
    n = L.ref_write_colors(rgb.ctypes.data, W, H, None, 0)
    buf = np.empty(n, np.uint8)
    L.ref_write_colors(rgb.ctypes.data, W, H, buf.ctypes.data, n)
    return buf.tobytes()


def _reflatten(self, float_images=False, bilinear=False):
    """The same live reference objects flattened with the adaptor's texture opt-ins -> Scene (the handle's description changes too)."""
    L = lib()
    L.ref_scene_reflatten.argtypes = [C.c_int, C.c_int, C.c_int]
    if L.ref_scene_reflatten(self.h, int(float_images), int(bilinear)) != 0:
        raise RuntimeError("ref_scene_reflatten: " + L.ref_last_error().decode())
    sc = Scene.from_desc_pointer(L.ref_scene_desc(self.h))
    sc.name = self.name
    return sc


RefScene.reflatten = _reflatten
