"""Host-side mirror of the reference camera initialisers (src/camera.h:21-132).

Same names, argument order, defaults and float/double mix as the reference, producing the public camera
fields the C ABI takes (b200rt_camera).  `image_height = int(image_width / aspect_ratio)` is evaluated in
double for perspective / orthnormal / fisheye and in FLOAT for lens (its aspect_ratio parameter is a float,
src/camera.h:104); theta is a float everywhere (:45).
"""
import math

import numpy as np

from . import _abi as abi

_PI = 3.1415926535897932385  # src/utility.h:15


def _degrees_to_radians(deg):  # src/utility.h: degrees * pi / 180.0
    return deg * _PI / 180.0


def _unit(v):
    return v / math.sqrt(v[0] * v[0] + v[1] * v[1] + v[2] * v[2])


def _cross(a, b):
    return np.array([a[1] * b[2] - a[2] * b[1], a[2] * b[0] - a[0] * b[2], a[0] * b[1] - a[1] * b[0]])


def _frame(cam, pos, lookat):
    pos = np.asarray(pos, dtype=np.float64)
    lookat = np.asarray(lookat, dtype=np.float64)
    d = _unit(lookat - pos)
    right = _unit(_cross(d, np.array([0.0, 1.0, 0.0])))
    up = _cross(right, d)
    for k in range(3):
        cam.pos[k], cam.dir[k], cam.right[k], cam.up[k] = pos[k], d[k], right[k], up[k]
    return right, up


def _size(cam, image_width, aspect_ratio, spp, depth):
    cam.image_width = int(image_width)
    h = int(image_width / aspect_ratio)
    cam.image_height = 1 if h < 1 else h
    cam.samples_per_pixel = int(spp)
    cam.max_depth = int(depth)


def _viewport(cam, fovy_degree, length):
    theta = float(np.float32(_degrees_to_radians(float(np.float32(fovy_degree)))))
    cam.viewport_height = 2.0 * math.tan(theta / 2.0) * float(np.float32(length))
    cam.viewport_width = cam.viewport_height * (float(cam.image_width) / cam.image_height)


def initialize_perspective(image_width, aspect_ratio, pos, lookat, focal_length=1, fovy_degree=90, sample_per_pixel=100, max_recur_depth=5):
    cam = abi.Camera()
    cam.mode = abi.CAM_PERSPECTIVE
    _frame(cam, pos, lookat)
    _size(cam, image_width, float(aspect_ratio), sample_per_pixel, max_recur_depth)
    cam.focal_length = float(np.float32(focal_length))
    cam.focus_dist = 3.4  # member default (src/camera.h:315); only the lens camera reads it
    _viewport(cam, fovy_degree, focal_length)
    return cam


def initialize_orthnormal(image_width, aspect_ratio, viewport_height, pos, lookat, sample_per_pixel=100, max_recur_depth=5):
    cam = abi.Camera()
    cam.mode = abi.CAM_ORTHOGRAPHIC
    _frame(cam, pos, lookat)
    _size(cam, image_width, float(aspect_ratio), sample_per_pixel, max_recur_depth)
    cam.focal_length = 0.0  # uninitialised in the reference and never read by the orthographic generate_ray (:252-258)
    cam.focus_dist = 3.4
    cam.viewport_height = float(viewport_height)
    cam.viewport_width = float(viewport_height) * (float(cam.image_width) / cam.image_height)
    return cam


def initialize_fisheye(image_width, aspect_ratio, pos, lookat, focal_length=1, fovy_degree=90, sample_per_pixel=100, max_recur_depth=5):
    cam = initialize_perspective(image_width, aspect_ratio, pos, lookat, focal_length, fovy_degree, sample_per_pixel, max_recur_depth)
    cam.mode = abi.CAM_FISHEYE
    return cam


def initialize_lens(image_width, aspect_ratio, pos, lookat, defocus_angle, focus_dist=1, fovy_degree=90, sample_per_pixel=100, max_recur_depth=5):
    cam = abi.Camera()
    cam.mode = abi.CAM_LENS
    right, up = _frame(cam, pos, lookat)
    cam.image_width = int(image_width)
    h = int(np.float32(image_width) / np.float32(aspect_ratio))  # float division: aspect_ratio is a float here
    cam.image_height = 1 if h < 1 else h
    cam.samples_per_pixel = int(sample_per_pixel)
    cam.max_depth = int(max_recur_depth)
    fd = float(np.float32(focus_dist))
    cam.focus_dist = fd
    cam.focal_length = 0.0  # never read in lens mode (src/camera.h:276-283)
    _viewport(cam, fovy_degree, focus_dist)
    # focus_dist * std::tan(degrees_to_radians(defocus_angle_ / 2)): float / int -> float, then double math, result float*double
    half = float(np.float32(defocus_angle) / np.float32(2))
    radius = fd * math.tan(_degrees_to_radians(half))
    for k in range(3):
        cam.defocus_disk_u[k] = right[k] * radius
        cam.defocus_disk_v[k] = up[k] * radius
    return cam
