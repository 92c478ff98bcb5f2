"""ctypes front end of tests/emu/libemu.so — TEST INFRASTRUCTURE (see emu.cc)."""
import ctypes as C
import importlib
import os
import subprocess
import sys

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_ROOT = os.path.dirname(os.path.dirname(_HERE))
if _ROOT not in sys.path:
    sys.path.insert(0, _ROOT)
_pkg = importlib.import_module("cpu-ray-tracing-implementation_b200")
abi = _pkg._abi
LIB = os.path.join(_HERE, "libemu.so")
CSRC = os.path.join(_ROOT, "cpu-ray-tracing-implementation_b200", "csrc")


def build(force=False):
    srcs = [os.path.join(_HERE, "emu.cc"), os.path.join(_HERE, "host_emu.h")] + [os.path.join(CSRC, f) for f in os.listdir(CSRC)]
    if not force and os.path.exists(LIB) and all(os.path.getmtime(LIB) >= os.path.getmtime(s) for s in srcs):
        return LIB
    cmd = ["g++", "-O2", "-std=c++17", "-ffp-contract=off", "-fPIC", "-shared", "-DB200_HOST_EMU", "-I" + _HERE, "-I" + CSRC,
           "-I" + os.path.join(_ROOT, "include"), os.path.join(_HERE, "emu.cc"), "-o", LIB]
    subprocess.run(cmd, check=True)
    return LIB


_lib = None


def lib():
    global _lib
    if _lib is None:
        L = C.CDLL(build())
        vp, i32, u64, dbl, sz = C.c_void_p, C.c_int, C.c_uint64, C.c_double, C.c_size_t
        L.emu_scene_create.argtypes = [C.POINTER(abi.SceneDesc), i32]
        L.emu_scene_create.restype = vp
        L.emu_scene_destroy.argtypes = [vp]
        L.emu_scene_info.argtypes = [vp, vp]
        L.emu_trace.argtypes = [vp, i32, vp, sz, dbl, dbl, u64, vp, vp, vp, vp, vp, vp]
        L.emu_generate_rays.argtypes = [C.POINTER(abi.Camera), u64, i32, i32, vp]
        L.emu_texture_sample.argtypes = [vp, i32, vp, sz, vp]
        L.emu_scatter.argtypes = [vp, vp, vp, vp, sz, vp]
        L.emu_render.argtypes = [vp, i32, C.POINTER(abi.Camera), u64, i32, i32, i32, vp, vp, vp]
        _lib = L
    return _lib


class EmuScene:
    def __init__(self, scene, leaf_size=2):
        self.scene = scene
        self._d, self._keep = scene.desc()
        self.h = lib().emu_scene_create(C.byref(self._d), leaf_size)
        if not self.h:
            raise RuntimeError("emu_scene_create failed")

    def bvh_hash(self):
        f = lib().emu_scene_bvh_hash
        f.argtypes, f.restype = [C.c_void_p], C.c_uint64
        return f(self.h)

    def info(self):
        a = np.zeros(6, np.uint32)
        lib().emu_scene_info(self.h, a.ctypes.data)
        return dict(zip(("bvh_nodes", "leaves", "entries", "bvh_depth", "graph_depth", "chains"), a.tolist()))

    def trace(self, rays, tmin=0.001, tmax=np.inf, seed=1, mirror=False):
        rays = np.ascontiguousarray(rays, dtype=np.float64).reshape(-1, 7)
        n = len(rays)
        out = dict(prim=np.zeros(n, np.int32), t=np.zeros(n), uv=np.zeros((n, 2)), front=np.zeros(n, np.uint8), p=np.zeros((n, 3)),
                   normal=np.zeros((n, 3)))
        lib().emu_trace(self.h, int(mirror), rays.ctypes.data, n, tmin, tmax, seed, *(out[k].ctypes.data for k in ("prim", "t", "uv", "front", "p", "normal")))
        return out

    def texture_sample(self, tex, uvp):
        uvp = np.ascontiguousarray(uvp, dtype=np.float64).reshape(-1, 5)
        rgb = np.zeros((len(uvp), 3))
        lib().emu_texture_sample(self.h, tex, uvp.ctypes.data, len(uvp), rgb.ctypes.data)
        return rgb

    def scatter(self, rays, rec12, draws6):
        rays = np.ascontiguousarray(rays, dtype=np.float64).reshape(-1, 7)
        rec12 = np.ascontiguousarray(rec12, dtype=np.float64).reshape(-1, 12)
        draws6 = np.ascontiguousarray(draws6, dtype=np.float64).reshape(-1, 6)
        out = np.zeros((len(rays), 16))
        lib().emu_scatter(self.h, rays.ctypes.data, rec12.ctypes.data, draws6.ctypes.data, len(rays), out.ctypes.data)
        return out

    def render(self, camera, seed, spp_begin, spp_end, mirror=False, kill_zero=False):
        H, W = camera.image_height, camera.image_width
        s = np.zeros((H, W, 3))
        ps = np.zeros((spp_end - spp_begin, H, W, 3))
        rays = C.c_uint64(0)
        lib().emu_render(self.h, int(mirror), C.byref(camera), seed, spp_begin, spp_end, int(kill_zero), s.ctypes.data, ps.ctypes.data, C.byref(rays))
        return dict(sum=s, samples=ps, rays=rays.value)

    def __del__(self):
        try:
            lib().emu_scene_destroy(self.h)
        except Exception:
            pass


def generate_rays(camera, seed, spp_begin, spp_end):
    n = (spp_end - spp_begin) * camera.image_width * camera.image_height
    rays = np.zeros((n, 7))
    lib().emu_generate_rays(C.byref(camera), seed, spp_begin, spp_end, rays.ctypes.data)
    return rays
