// camera.cuh — camera::generate_ray (src/camera.h:244-284) for the four camera modes, on Philox draws.
#pragma once
#include "scene_dev.cuh"

namespace b200 {

struct DevCamera {
  b200rt_camera c;
  V delta_u, delta_v; // src/camera.h:137-138
  V d00;              // perspective / fisheye: direction through the centre of pixel (0,0)   (:246)
  V p00;              // orthographic / lens: centre of pixel (0,0) on the viewport plane     (:253, :277)
};

HD DevCamera make_camera(const b200rt_camera &c) {
  DevCamera k;
  k.c = c;
  const V pos = v3(c.pos[0], c.pos[1], c.pos[2]), dir = v3(c.dir[0], c.dir[1], c.dir[2]);
  const V right = v3(c.right[0], c.right[1], c.right[2]), up = v3(c.up[0], c.up[1], c.up[2]);
  k.delta_u = c.viewport_width * right / c.image_width;
  k.delta_v = -c.viewport_height * up / c.image_height;
  k.d00 = c.focal_length * dir - c.viewport_width / 2.0 * right + c.viewport_height / 2.0 * up + 0.5 * (k.delta_u + k.delta_v);
  k.p00 = pos - c.viewport_width / 2.0 * right + c.viewport_height / 2.0 * up + 0.5 * (k.delta_u + k.delta_v);
  return k;
}

// y = row (i), x = column (j).  Draw use: P_CAMERA {u0,u1 = sample_square (:293), u2 = ray time};
// P_LENS aux=it {u0,u1} = the it-th rejection round of random_in_unit_disk (src/utility.h:46-52).
// FAST_DEAD (the render path): a fisheye pixel whose whole sample square lies outside the image circle returns its NaN ray without
// drawing (its time, which only b200rt_generate_rays reports, is then 0 instead of the draw).
template <bool FAST_DEAD = false>
DEV Ray generate_ray(const DevCamera &k, const Stream &st, int y, int x) {
  const b200rt_camera &c = k.c;
  if (FAST_DEAD && c.mode == B200RT_CAM_FISHEYE) {
    const V dir0 = v3(c.dir[0], c.dir[1], c.dir[2]);
    const V rd0 = k.d00 + x * k.delta_u + y * k.delta_v;
    // |offset| <= 0.5 (|delta_u| + |delta_v|) for any sample of the pixel, so rr >= |rd0 - dir| - that; 1e-6 covers every rounding
    const double least = length(rd0 - dir0) - 0.5 * (length(k.delta_u) + length(k.delta_v));
    if (least > c.focal_length * (1.0 + 1e-6)) {
      const double qnan = asin(2.0);
      return Ray{v3(c.pos[0], c.pos[1], c.pos[2]), v3(qnan, qnan, qnan), 0.0};
    }
  }
  const Draws dr = st.get(0, P_CAMERA, 0);
  const double ox = dr.u0 - 0.5, oy = dr.u1 - 0.5;
  const V pos = v3(c.pos[0], c.pos[1], c.pos[2]), dir = v3(c.dir[0], c.dir[1], c.dir[2]);
  if (c.mode == B200RT_CAM_PERSPECTIVE) { // :245-251, direction not normalised
    const V rd = k.d00 + x * k.delta_u + y * k.delta_v;
    return Ray{pos, rd + ox * k.delta_u + oy * k.delta_v, dr.u2};
  } else if (c.mode == B200RT_CAM_ORTHOGRAPHIC) { // :252-258
    const V p = k.p00 + x * k.delta_u + y * k.delta_v;
    return Ray{p + ox * k.delta_u + oy * k.delta_v, dir, dr.u2};
  } else if (c.mode == B200RT_CAM_FISHEYE) { // :259-275; NaN outside the image circle (asin of > 1)
    const V rd = k.d00 + x * k.delta_u + y * k.delta_v;
    const V rrd = rd + ox * k.delta_u + oy * k.delta_v;
    const double rr = length(rrd - dir);
    const double sin_arg = rr / c.focal_length;
    // outside the image circle asin() is NaN and every later product with it too: the reference's ray is NaN in all three direction
    // components.  Return that ray before the asin / sin / cos / sqrt chain: 56 % of the Fox config's samples
    if (!(sin_arg <= 1.0)) {
      const double qnan = asin(2.0);
      return Ray{pos, v3(qnan, qnan, qnan), dr.u2};
    }
    const double theta = asin(sin_arg);
    const V v1 = unit_vector(dir), v2 = unit_vector(rrd);
    const double sn = sin(theta);
    const double b_prime = sqrt(sn * sn / (1 - dot(v1, v2) * dot(v1, v2)));
    const double a_prime = cos(theta) - b_prime * dot(v1, v2);
    return Ray{pos, a_prime * v1 + b_prime * v2, dr.u2};
  } else { // lens, :276-283: time = 0
    V rd = k.p00 + x * k.delta_u + y * k.delta_v + ox * k.delta_u + oy * k.delta_v + c.focus_dist * dir;
    V p;
    for (uint32_t it = 0;; it++) {
      const Draws dl = st.get(0, P_LENS, it);
      p = v3(-1 + (1 - -1) * dl.u0, -1 + (1 - -1) * dl.u1, 0);
      if (length_squared(p) < 1) break;
    }
    const V du = v3(c.defocus_disk_u[0], c.defocus_disk_u[1], c.defocus_disk_u[2]);
    const V dv = v3(c.defocus_disk_v[0], c.defocus_disk_v[1], c.defocus_disk_v[2]);
    const V ro = pos + ((p.x * du) + (p.y * dv));
    return Ray{ro, rd - ro, 0.0};
  }
}

} // namespace b200
