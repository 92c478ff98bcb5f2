"""ctypes front end of oracle/liboracle.so (oracle.cc, the CPU restatement) — TEST INFRASTRUCTURE.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may import this module.
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

LIB_PATH = os.path.join(_HERE, "liboracle.so")


class OracleStats(C.Structure):
    _fields_ = [("samples", C.c_uint64), ("rays", C.c_uint64), ("nan_samples", C.c_uint64), ("seconds", C.c_double)]


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(f"{LIB_PATH} missing: run `make -f oracle/Makefile` (or __graft_entry__.build())")
        L = C.CDLL(LIB_PATH)
        vp, i32, u64, dbl, sz = C.c_void_p, C.c_int, C.c_uint64, C.c_double, C.c_size_t
        SD, CAM = C.POINTER(abi.SceneDesc), C.POINTER(abi.Camera)
        L.oracle_trace.argtypes = [SD, vp, sz, dbl, dbl, u64, vp, vp, vp, vp, vp, vp]
        L.oracle_generate_rays.argtypes = [CAM, u64, i32, i32, vp]
        L.oracle_generate_rays_from_draws.argtypes = [CAM, sz, vp, i32, i32, vp]
        L.oracle_texture_sample.argtypes = [SD, i32, vp, sz, vp]
        L.oracle_scatter.argtypes = [SD, vp, vp, vp, sz, vp]
        L.oracle_ray_color.argtypes = [SD, vp, sz, i32, u64, vp]
        L.oracle_render.argtypes = [SD, CAM, u64, i32, i32, i32, vp, vp, vp, C.POINTER(OracleStats)]
        _lib = L
    return _lib


def trace(scene, rays, tmin=0.001, tmax=np.inf, seed=1):
    d, keep = scene.desc()
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
    rc = lib().oracle_trace(C.byref(d), rays.ctypes.data, n, tmin, tmax, seed, *(out[k].ctypes.data for k in ("prim", "t", "uv", "front", "p", "normal")))
    assert rc == 0
    return out


def generate_rays(camera, seed, spp_begin, spp_end):
    n = (spp_end - spp_begin) * camera.image_width * camera.image_height
    rays = np.zeros((n, 7), np.float64)
    assert lib().oracle_generate_rays(C.byref(camera), seed, spp_begin, spp_end, rays.ctypes.data) == 0
    return rays


def generate_rays_from_draws(camera, draws, swap_pairs=True):
    draws = np.ascontiguousarray(draws, dtype=np.float64)
    n, m = draws.shape
    rays = np.zeros((n, 7), np.float64)
    assert lib().oracle_generate_rays_from_draws(C.byref(camera), n, draws.ctypes.data, m, int(swap_pairs), rays.ctypes.data) == 0
    return rays


def texture_sample(scene, tex, uvp):
    d, keep = scene.desc()
    uvp = np.ascontiguousarray(uvp, dtype=np.float64).reshape(-1, 5)
    rgb = np.zeros((len(uvp), 3), np.float64)
    assert lib().oracle_texture_sample(C.byref(d), tex, uvp.ctypes.data, len(uvp), rgb.ctypes.data) == 0
    return rgb


def scatter(scene, rays, rec12, draws6):
    d, keep = scene.desc()
    rays = np.ascontiguousarray(rays, dtype=np.float64).reshape(-1, 7)
    rec12 = np.ascontiguousarray(rec12, dtype=np.float64).reshape(-1, 12)
    draws6 = np.ascontiguousarray(draws6, dtype=np.float64).reshape(-1, 6)
    out = np.zeros((len(rays), 16), np.float64)
    assert lib().oracle_scatter(C.byref(d), rays.ctypes.data, rec12.ctypes.data, draws6.ctypes.data, len(rays), out.ctypes.data) == 0
    return out


def ray_color(scene, rays, depth, seed=1):
    d, keep = scene.desc()
    rays = np.ascontiguousarray(rays, dtype=np.float64).reshape(-1, 7)
    rgb = np.zeros((len(rays), 3), np.float64)
    assert lib().oracle_ray_color(C.byref(d), rays.ctypes.data, len(rays), depth, seed, rgb.ctypes.data) == 0
    return rgb


def render(scene, camera, seed, spp_begin, spp_end, threads=0, want_sumsq=False, want_samples=False):
    """-> dict(sum=(H,W,3), sumsq=..., samples=(spp,H,W,3), stats=dict)"""
    d, keep = scene.desc()
    if threads <= 0:
        threads = os.cpu_count() or 1
    H, W = camera.image_height, camera.image_width
    s = np.zeros((H, W, 3), np.float64)
    q = np.zeros((H, W, 3), np.float64) if want_sumsq else None
    ps = np.zeros((spp_end - spp_begin, H, W, 3), np.float64) if want_samples else None
    st = OracleStats()
    rc = lib().oracle_render(C.byref(d), C.byref(camera), seed, spp_begin, spp_end, threads, s.ctypes.data,
                             q.ctypes.data if want_sumsq else None, ps.ctypes.data if want_samples else None, C.byref(st))
    assert rc == 0
    return dict(sum=s, sumsq=q, samples=ps, stats={k: getattr(st, k) for k, _ in OracleStats._fields_})
