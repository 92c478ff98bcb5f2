"""ctypes / numpy mirror of include/b200rt.h (the C ABI of libb200rt.so).

Every structure here is laid out exactly like its C counterpart; tests/test_abi.py checks the
sizes against the compiled library (b200rt_abi_sizeof_*), so drift fails loudly.
"""
import ctypes as C

import numpy as np

ABI_VERSION = 2

# ---- enums (include/b200rt.h) -------------------------------------------------------------------
NODE_BVH, NODE_LIST, NODE_TRANSLATE, NODE_ROTATE_X, NODE_ROTATE_Y, NODE_ROTATE_Z = 0, 1, 2, 3, 4, 5
NODE_SPHERE, NODE_QUAD, NODE_TRIANGLE, NODE_VOLUME = 6, 7, 8, 9
MAT_LAMBERTIAN, MAT_METAL, MAT_DIELECTRIC, MAT_GLOSS, MAT_ISOTROPIC, MAT_DIFFUSE_LIGHT = 0, 1, 2, 3, 4, 5
TEX_SOLID, TEX_CHECKER, TEX_PICTURE, TEX_PERLIN, TEX_VALUE, TEX_WORLEY, TEX_VORONOI, TEX_PICTURE_LINEAR = 0, 1, 2, 3, 4, 5, 6, 7
LIGHT_NONE, LIGHT_QUAD, LIGHT_SPHERE, LIGHT_DEFAULT = 0, 1, 2, 3
CAM_PERSPECTIVE, CAM_ORTHOGRAPHIC, CAM_FISHEYE, CAM_LENS = 0, 1, 2, 3
OPT_TRAVERSAL, OPT_WAVE_PATHS, OPT_KILL_ZERO_WEIGHT, OPT_BVH_LEAF_SIZE, OPT_RESERVED_4, OPT_REFILL_LANES, OPT_EXTEND_BLOCKS, OPT_SPECULATIVE_LEAF, OPT_FUSE_MISS, OPT_EXTEND_STAGES, OPT_SMEM_CARVEOUT, OPT_BVH_BUILDER, OPT_BVH_WIDTH, OPT_PIPELINES, OPT_PIPE_EXTEND_CTAS, OPT_TAIL_PATHS, OPT_WAVE_LOOP, OPT_EXTEND_LOOP, OPT_LEAF_LANES, OPT_L2_PERSIST, OPT_TRI_F32, OPT_OCTANT_SORT, OPT_RAY_SORT = 0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11, 12, 13, 14, 15, 16, 17, 18, 19, 20, 21, 22
BVH_HOST_SAH, BVH_DEVICE_LBVH, BVH_DEVICE_PLOC = 0, 1, 2
TRAVERSAL_FAST, TRAVERSAL_MIRROR = 0, 1

SPHERE_STRIDE, QUAD_STRIDE, TRI_STRIDE = 12, 20, 12

# ---- record dtypes (numpy, C layout) ------------------------------------------------------------
node_dtype = np.dtype([("kind", "<i4"), ("a", "<i4"), ("b", "<i4"), ("c", "<i4"), ("f", "<f8", (6,))], align=True)
volume_dtype = np.dtype([("boundary_node", "<i4"), ("material", "<i4"), ("density", "<f8")], align=True)
material_dtype = np.dtype([("kind", "<i4"), ("tex", "<i4"), ("p0", "<f4"), ("p1", "<f4")], align=True)
texture_dtype = np.dtype(
    [("kind", "<i4"), ("aux", "<i4"), ("c0", "<f8", (3,)), ("c1", "<f8", (3,)), ("scale", "<f8")], align=True
)
image_dtype = np.dtype(
    [("width", "<i4"), ("height", "<i4"), ("bytes_per_pixel", "<i4"), ("has_data", "<i4"), ("offset", "<u8")], align=True
)
value_noise_dtype = np.dtype([("resolution", "<i4"), ("pad", "<i4"), ("offset", "<u8")], align=True)

assert node_dtype.itemsize == 64 and volume_dtype.itemsize == 16 and material_dtype.itemsize == 16
assert texture_dtype.itemsize == 64 and image_dtype.itemsize == 24 and value_noise_dtype.itemsize == 16


class SceneDesc(C.Structure):
    _fields_ = [
        ("abi_version", C.c_uint32),
        ("root", C.c_uint32),
        ("n_nodes", C.c_uint32),
        ("n_list_children", C.c_uint32),
        ("nodes", C.c_void_p),
        ("list_children", C.c_void_p),
        ("n_spheres", C.c_uint32),
        ("n_quads", C.c_uint32),
        ("n_tris", C.c_uint32),
        ("n_volumes", C.c_uint32),
        ("spheres", C.c_void_p),
        ("sphere_mat", C.c_void_p),
        ("quads", C.c_void_p),
        ("quad_mat", C.c_void_p),
        ("tris", C.c_void_p),
        ("tri_mat", C.c_void_p),
        ("volumes", C.c_void_p),
        ("n_materials", C.c_uint32),
        ("n_textures", C.c_uint32),
        ("materials", C.c_void_p),
        ("textures", C.c_void_p),
        ("n_images", C.c_uint32),
        ("n_perlin", C.c_uint32),
        ("n_value_noise", C.c_uint32),
        ("n_primitives", C.c_uint32),
        ("images", C.c_void_p),
        ("image_bytes", C.c_void_p),
        ("image_bytes_size", C.c_uint64),
        ("perlin_vec", C.c_void_p),
        ("perlin_perm", C.c_void_p),
        ("value_noise", C.c_void_p),
        ("value_lattice", C.c_void_p),
        ("value_lattice_size", C.c_uint64),
        ("light_kind", C.c_int32),
        ("light_index", C.c_int32),
        ("background_tex", C.c_int32),
        ("reserved", C.c_int32),
        ("tri_uv", C.c_void_p),
    ]


class Camera(C.Structure):
    _fields_ = [
        ("mode", C.c_int32),
        ("image_width", C.c_int32),
        ("image_height", C.c_int32),
        ("samples_per_pixel", C.c_int32),
        ("max_depth", C.c_int32),
        ("reserved", C.c_int32),
        ("viewport_width", C.c_double),
        ("viewport_height", C.c_double),
        ("focal_length", C.c_double),
        ("focus_dist", C.c_double),
        ("pos", C.c_double * 3),
        ("dir", C.c_double * 3),
        ("right", C.c_double * 3),
        ("up", C.c_double * 3),
        ("defocus_disk_u", C.c_double * 3),
        ("defocus_disk_v", C.c_double * 3),
    ]

    def copy(self):
        c = Camera()
        C.memmove(C.byref(c), C.byref(self), C.sizeof(Camera))
        return c

    def to_bytes(self):
        return bytes(C.string_at(C.byref(self), C.sizeof(Camera)))

    @classmethod
    def from_bytes(cls, b):
        c = cls()
        C.memmove(C.byref(c), bytes(b), C.sizeof(cls))
        return c


class Stats(C.Structure):
    _fields_ = [
        ("samples", C.c_uint64),
        ("rays", C.c_uint64),
        ("nan_samples", C.c_uint64),
        ("kernel_launches", C.c_uint64),
        ("gpu_ms", C.c_double),
        ("extend_ms", C.c_double),
        ("h2d_bytes", C.c_double),
        ("d2h_bytes", C.c_double),
        ("waves", C.c_uint64),
    ]

    def as_dict(self):
        return {n: getattr(self, n) for n, _ in self._fields_}


class BvhInfo(C.Structure):
    _fields_ = [
        ("n_nodes", C.c_uint32),
        ("n_leaves", C.c_uint32),
        ("n_prims", C.c_uint32),
        ("max_depth", C.c_uint32),
        ("build_ms", C.c_double),
        ("sah_cost", C.c_double),
    ]


# Entry points every build of libb200rt.so must export (tests/test_abi.py walks this list and
# include/b200rt.h and checks they agree).
EXPORTS = [
    "b200rt_create",
    "b200rt_destroy",
    "b200rt_last_error",
    "b200rt_abi_version",
    "b200rt_set_option",
    "b200rt_upload_scene",
    "b200rt_set_camera",
    "b200rt_trace",
    "b200rt_generate_rays",
    "b200rt_render",
    "b200rt_render_device",
    "b200rt_finalize",
    "b200rt_ppm_bound",
    "b200rt_format_ppm",
    "b200rt_bvh_info_get",
    "b200rt_texture_sample",
    "b200rt_scatter",
    "b200rt_abi_sizeof",
    "b200rt_image_bound",
    "b200rt_format_image",
    "b200rt_format_ppm_noalloc",
    "b200rt_group_create",
    "b200rt_group_destroy",
    "b200rt_group_size",
    "b200rt_group_context",
    "b200rt_group_set_option",
    "b200rt_group_upload_scene",
    "b200rt_group_set_camera",
    "b200rt_group_render",
    "b200rt_group_reduce_backend",
]
GROUP_OPT_REDUCE, GROUP_REDUCE_NCCL, GROUP_REDUCE_PEER = 1000, 0, 1
IMAGE_PPM_TEXT, IMAGE_PPM_BINARY, IMAGE_EXR = 0, 1, 2
NOALLOC_MAX_PIXELS = 1 << 20  # B200RT_NOALLOC_MAX_PIXELS: b200rt_format_ppm_noalloc's static scratch


class _Missing:
    """Stands in for the symbols an OLDER build of the library lacks (only when B200RT_LIB names such a build for an A/B run)."""

    def __getattr__(self, name):
        return self

    def __setattr__(self, name, value):
        pass


def bind(lib):
    """Attach argtypes/restypes of include/b200rt.h to a loaded libb200rt.so."""
    import os

    if os.environ.get("B200RT_LIB"):
        real = lib

        class _Tolerant:
            def __getattr__(self, name):
                try:
                    return getattr(real, name)
                except AttributeError:
                    return _Missing()

        _bind(_Tolerant())
        return real
    return _bind(lib)


def _bind(lib):
    vp, i32, u64, dbl, sz = C.c_void_p, C.c_int, C.c_uint64, C.c_double, C.c_size_t
    lib.b200rt_create.argtypes = [i32, C.POINTER(vp)]
    lib.b200rt_destroy.argtypes = [vp]
    lib.b200rt_last_error.restype = C.c_char_p
    lib.b200rt_last_error.argtypes = []
    lib.b200rt_abi_version.argtypes = []
    lib.b200rt_set_option.argtypes = [vp, i32, dbl]
    lib.b200rt_upload_scene.argtypes = [vp, C.POINTER(SceneDesc)]
    lib.b200rt_set_camera.argtypes = [vp, C.POINTER(Camera)]
    lib.b200rt_trace.argtypes = [vp, vp, sz, dbl, dbl, u64, vp, vp, vp, vp, vp, vp]
    lib.b200rt_generate_rays.argtypes = [vp, u64, i32, i32, vp]
    lib.b200rt_render.argtypes = [vp, u64, i32, i32, vp, C.POINTER(Stats)]
    lib.b200rt_render_device.argtypes = [vp, u64, i32, i32, vp, vp, C.POINTER(Stats)]
    lib.b200rt_finalize.argtypes = [vp, vp, i32, vp, vp]
    lib.b200rt_ppm_bound.argtypes = [i32, i32]
    lib.b200rt_format_ppm.argtypes = [vp, vp, i32, vp, sz, C.POINTER(sz)]
    lib.b200rt_bvh_info_get.argtypes = [vp, C.POINTER(BvhInfo)]
    lib.b200rt_texture_sample.argtypes = [vp, i32, vp, sz, vp]
    lib.b200rt_scatter.argtypes = [vp, vp, vp, vp, sz, vp]
    lib.b200rt_abi_sizeof.argtypes = [i32]
    lib.b200rt_image_bound.argtypes = [i32, i32, i32]
    lib.b200rt_image_bound.restype = C.c_size_t
    lib.b200rt_format_image.argtypes = [vp, vp, i32, i32, vp, sz, C.POINTER(sz)]
    lib.b200rt_format_ppm_noalloc.argtypes = [vp, vp, i32, vp, sz, C.POINTER(sz)]  # device pointers in and out
    lib.b200rt_group_create.argtypes = [C.POINTER(i32), i32, C.POINTER(vp)]
    lib.b200rt_group_destroy.argtypes = [vp]
    lib.b200rt_group_size.argtypes = [vp]
    lib.b200rt_group_context.argtypes = [vp, i32, C.POINTER(vp)]
    lib.b200rt_group_set_option.argtypes = [vp, i32, dbl]
    lib.b200rt_group_upload_scene.argtypes = [vp, C.POINTER(SceneDesc)]
    lib.b200rt_group_set_camera.argtypes = [vp, C.POINTER(Camera)]
    lib.b200rt_group_render.argtypes = [vp, u64, i32, i32, vp, C.POINTER(Stats)]
    lib.b200rt_group_reduce_backend.argtypes = []
    lib.b200rt_group_reduce_backend.restype = C.c_char_p
    for name in EXPORTS:
        if name not in ("b200rt_last_error", "b200rt_abi_sizeof", "b200rt_ppm_bound", "b200rt_group_reduce_backend", "b200rt_image_bound"):
            getattr(lib, name).restype = C.c_int
    lib.b200rt_abi_sizeof.restype = C.c_size_t
    lib.b200rt_ppm_bound.restype = C.c_size_t
    return lib
