"""ctypes binding of libb200rt.so — the product path.

There is no CPU fallback: if the CUDA library is missing or fails to load, importing the
binding raises; if a call fails, `B200RTError` carries b200rt_last_error().
"""
import ctypes as C
import os

import numpy as np

from . import _abi as abi

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("B200RT_LIB") or os.path.join(_HERE, "libb200rt.so")  # B200RT_LIB: another build of the SAME library (A/B experiments)


class B200RTError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"b200rt error {code}: {msg}")
        self.code = code


_lib = None


def lib():
    """Load libb200rt.so (built in-tree by __graft_entry__.build()).  Fails loudly when absent."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'`; "
                               "this package has no CPU fallback")
        _lib = abi.bind(C.CDLL(LIB_PATH))
    return _lib


def _ptr(a):
    return a.ctypes.data if a is not None else None


class Context:
    """One b200rt context = one GPU."""

    def __init__(self, device=0):
        self._h = C.c_void_p()
        self._lib = lib()
        self._check(self._lib.b200rt_create(device, C.byref(self._h)))
        self.device = device
        self.camera = None
        self._keep = None

    def _check(self, rc):
        if rc != 0:
            raise B200RTError(rc, self._lib.b200rt_last_error().decode())

    def close(self):
        if self._h:
            self._lib.b200rt_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    # ---- configuration
    def set_option(self, option, value):
        self._check(self._lib.b200rt_set_option(self._h, option, float(value)))

    def upload_scene(self, scene):
        d, keep = scene.desc()
        self._check(self._lib.b200rt_upload_scene(self._h, C.byref(d)))

    def set_camera(self, camera):
        self.camera = camera.copy()
        self._check(self._lib.b200rt_set_camera(self._h, C.byref(self.camera)))

    def bvh_info(self):
        info = abi.BvhInfo()
        self._check(self._lib.b200rt_bvh_info_get(self._h, C.byref(info)))
        return {n: getattr(info, n) for n, _ in abi.BvhInfo._fields_}

    # ---- batch entry points
    def trace(self, rays, tmin=0.001, tmax=np.inf, seed=1, want=("prim", "t", "uv", "front", "p", "normal")):
        rays = np.ascontiguousarray(rays, dtype=np.float64).reshape(-1, 7)
        n = len(rays)
        shapes = dict(prim=((n,), np.int32), t=((n,), np.float64), uv=((n, 2), np.float64), front=((n,), np.uint8),
                      p=((n, 3), np.float64), normal=((n, 3), np.float64))
        out = {k: np.zeros(*shapes[k]) for k in want}
        args = [_ptr(out.get(k)) for k in ("prim", "t", "uv", "front", "p", "normal")]
        self._check(self._lib.b200rt_trace(self._h, rays.ctypes.data, n, tmin, tmax, seed, *args))
        return out

    def texture_sample(self, tex, uvp):
        uvp = np.ascontiguousarray(uvp, dtype=np.float64).reshape(-1, 5)
        rgb = np.zeros((len(uvp), 3), np.float64)
        self._check(self._lib.b200rt_texture_sample(self._h, tex, uvp.ctypes.data, len(uvp), rgb.ctypes.data))
        return rgb

    def scatter(self, rays, rec12, draws6):
        """One level of ray_color below the hit test on supplied hit records and supplied draws (b200rt_scatter) -> (n, 16)."""
        rays = np.ascontiguousarray(rays, dtype=np.float64).reshape(-1, 7)
        rec12 = np.ascontiguousarray(rec12, dtype=np.float64).reshape(-1, 12)
        draws6 = np.ascontiguousarray(draws6, dtype=np.float64).reshape(-1, 6)
        out = np.zeros((len(rays), 16), np.float64)
        self._check(self._lib.b200rt_scatter(self._h, rays.ctypes.data, rec12.ctypes.data, draws6.ctypes.data, len(rays), out.ctypes.data))
        return out

    def generate_rays(self, seed, spp_begin, spp_end):
        n = (spp_end - spp_begin) * self.camera.image_width * self.camera.image_height
        rays = np.zeros((n, 7), np.float64)
        self._check(self._lib.b200rt_generate_rays(self._h, seed, spp_begin, spp_end, rays.ctypes.data))
        return rays

    # ---- rendering
    def render(self, seed, spp_begin, spp_end, out=None, want_stats=True):
        """-> (rgb_sum (H, W, 3) float64 host array, stats dict | None)"""
        H, W = self.camera.image_height, self.camera.image_width
        if out is None:
            out = np.zeros((H, W, 3), np.float64)
        st = abi.Stats()
        self._check(self._lib.b200rt_render(self._h, seed, spp_begin, spp_end, out.ctypes.data, C.byref(st) if want_stats else None))
        return out, (st.as_dict() if want_stats else None)

    def render_device(self, seed, spp_begin, spp_end, data_ptr, stream=0, want_stats=False):
        """Accumulate into a device buffer of H*W*3 float64 (e.g. a torch tensor's data_ptr()) on CUDA stream `stream`."""
        st = abi.Stats()
        self._check(self._lib.b200rt_render_device(self._h, seed, spp_begin, spp_end, data_ptr, stream, C.byref(st) if want_stats else None))
        return st.as_dict() if want_stats else None

    def finalize(self, rgb_sum, total_spp, want_ppm=False):
        rgb_sum = np.ascontiguousarray(rgb_sum, dtype=np.float64)
        lin = np.zeros_like(rgb_sum)
        ppm = np.zeros(rgb_sum.shape, np.int32) if want_ppm else None
        self._check(self._lib.b200rt_finalize(self._h, rgb_sum.ctypes.data, total_spp, lin.ctypes.data, _ptr(ppm)))
        return (lin, ppm) if want_ppm else lin

    def format_ppm(self, rgb_sum, total_spp, capacity=None):
        """The complete P3 file image camera::render would write (src/camera.h:149-151, 174), as bytes; text made on the device."""
        rgb_sum = np.ascontiguousarray(rgb_sum, dtype=np.float64)
        cap = int(capacity if capacity is not None else self._lib.b200rt_ppm_bound(self.camera.image_width, self.camera.image_height))
        buf = np.empty(max(cap, 1), np.uint8)
        n = C.c_size_t(0)
        self._check(self._lib.b200rt_format_ppm(self._h, rgb_sum.ctypes.data, total_spp, buf.ctypes.data, cap, C.byref(n)))
        return buf[: n.value].tobytes()

    def format_ppm_noalloc(self, d_rgb_sum, total_spp, d_out, capacity):
        """b200rt_format_ppm_noalloc: DEVICE addresses in (W*H*3 doubles) and out (`capacity` bytes); -> the file's length.  The call
        allocates nothing: its scratch is a static device array for images of at most abi.NOALLOC_MAX_PIXELS pixels."""
        n = C.c_size_t(0)
        self._check(self._lib.b200rt_format_ppm_noalloc(self._h, int(d_rgb_sum), total_spp, int(d_out), int(capacity), C.byref(n)))
        return n.value

    def format_image(self, rgb_sum, total_spp, fmt, capacity=None):
        """The frame as a complete file image: abi.IMAGE_PPM_TEXT (the reference's P3), IMAGE_PPM_BINARY (P6) or IMAGE_EXR (linear float)."""
        rgb_sum = np.ascontiguousarray(rgb_sum, dtype=np.float64)
        cap = int(capacity if capacity is not None else self._lib.b200rt_image_bound(fmt, self.camera.image_width, self.camera.image_height))
        buf = np.empty(max(cap, 1), np.uint8)
        n = C.c_size_t(0)
        self._check(self._lib.b200rt_format_image(self._h, rgb_sum.ctypes.data, total_spp, fmt, buf.ctypes.data, cap, C.byref(n)))
        return buf[: n.value].tobytes()


class Group:
    """b200rt_group: the frame split by sample range across several GPUs of this box, one NCCL reduce (include/b200rt.h)."""

    def __init__(self, devices):
        self._lib = lib()
        self._h = C.c_void_p()
        arr = (C.c_int * len(devices))(*devices)
        self._check(self._lib.b200rt_group_create(arr, len(devices), C.byref(self._h)))
        self.devices = list(devices)
        self.camera = None

    def _check(self, rc):
        if rc != 0:
            raise B200RTError(rc, self._lib.b200rt_last_error().decode())

    def close(self):
        if self._h:
            self._lib.b200rt_group_destroy(self._h)
            self._h = C.c_void_p()

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_option(self, option, value):
        self._check(self._lib.b200rt_group_set_option(self._h, option, float(value)))

    def upload_scene(self, scene):
        d, keep = scene.desc()
        self._check(self._lib.b200rt_group_upload_scene(self._h, C.byref(d)))

    def set_camera(self, camera):
        self.camera = camera.copy()
        self._check(self._lib.b200rt_group_set_camera(self._h, C.byref(self.camera)))

    def render(self, seed, spp_begin, spp_end, out=None, want_stats=True):
        H, W = self.camera.image_height, self.camera.image_width
        if out is None:
            out = np.zeros((H, W, 3), np.float64)
        st = abi.Stats()
        self._check(self._lib.b200rt_group_render(self._h, seed, spp_begin, spp_end, out.ctypes.data, C.byref(st) if want_stats else None))
        return out, (st.as_dict() if want_stats else None)

    def reduce_backend(self):
        return self._lib.b200rt_group_reduce_backend().decode()
