// texture.cuh — texture::sample (src/texture.h:12-119) and the noise functions under it (src/noise.h).
// Lookup tables (perlin gradient vectors / permutation, value-noise lattice, 8-bit image rows) are the
// reference objects' own tables, copied at upload; nothing is regenerated on the device.
#pragma once
#include "scene_dev.cuh"

namespace b200 {

// perlin::perlin_interp, src/noise.h:22-43 (perm_x is used for all three axes: :36)
DEV double perlin_interp(const double *vec, const int32_t *perm, int iu, int iv, int iw, double u, double v, double w) {
  iu &= 255;
  iv &= 255;
  iw &= 255;
  const double uu = u * u * (3 - 2 * u), vv = v * v * (3 - 2 * v), ww = w * w * (3 - 2 * w);
  double accum = 0.0;
#pragma unroll
  for (int i = 0; i < 2; i++)
#pragma unroll
    for (int j = 0; j < 2; j++)
#pragma unroll
      for (int k = 0; k < 2; k++) {
        const V weight_v = v3(u - i, v - j, w - k);
        const int h = __ldg(perm + ((iu + i) & 255)) ^ __ldg(perm + ((iv + j) & 255)) ^ __ldg(perm + ((iw + k) & 255));
        const V g = v3(__ldg(vec + 3 * h), __ldg(vec + 3 * h + 1), __ldg(vec + 3 * h + 2));
        accum += (i * uu + (1 - i) * (1 - uu)) * (j * vv + (1 - j) * (1 - vv)) * (k * ww + (1 - k) * (1 - ww)) * dot(g, weight_v);
      }
  return accum;
}
DEV double perlin_noise(const double *vec, const int32_t *perm, V p) { // src/noise.h:57-69
  const int iu = int(floor(p.x)), iv = int(floor(p.y)), iw = int(floor(p.z));
  return perlin_interp(vec, perm, iu, iv, iw, p.x - iu, p.y - iv, p.z - iw);
}
DEV_COLD double perlin_turb(const double *vec, const int32_t *perm, int depth, V p) { // src/noise.h:45-55
  double accum = 0;
  V temp_p = p;
  double weight = 1.0;
  for (int i = 0; i < depth; i++) {
    accum += weight * perlin_noise(vec, perm, temp_p);
    weight *= 0.5;
    temp_p = temp_p * 2.0;
  }
  return fabs(accum);
}
DEV double lerp3(double t, double a, double b) { return (1 - t) * a + t * b; } // src/utility.h:84
DEV_COLD double value_noise(const float *lat, int res, V p) {                         // src/noise.h:106-132
  const V fp = vfloor(p);
  const double R = res;
  const float x0y0z0 = lat[(size_t)(fp.x * R * R + fp.y * R + fp.z)];
  const float x1y0z0 = lat[(size_t)((fp.x + 1) * R * R + fp.y * R + fp.z)];
  const float x0y1z0 = lat[(size_t)(fp.x * R * R + (fp.y + 1) * R + fp.z)];
  const float x1y1z0 = lat[(size_t)((fp.x + 1) * R * R + (fp.y + 1) * R + fp.z)];
  const float x0y0z1 = lat[(size_t)(fp.x * R * R + fp.y * R + fp.z + 1)];
  const float x1y0z1 = lat[(size_t)((fp.x + 1) * R * R + fp.y * R + fp.z + 1)];
  const float x0y1z1 = lat[(size_t)(fp.x * R * R + (fp.y + 1) * R + fp.z + 1)];
  const float x1y1z1 = lat[(size_t)((fp.x + 1) * R * R + (fp.y + 1) * R + fp.z + 1)];
  const double x = p.x - fp.x, y = p.y - fp.y, z = p.z - fp.z;
  const double y0z0 = lerp3(x, x0y0z0, x1y0z0), y1z0 = lerp3(x, x0y1z0, x1y1z0);
  const double y0z1 = lerp3(x, x0y0z1, x1y0z1), y1z1 = lerp3(x, x0y1z1, x1y1z1);
  const double z0 = lerp3(y, y0z0, y1z0), z1 = lerp3(y, y0z1, y1z1);
  return lerp3(z, z0, z1);
}
DEV V cell_offset(V u) { // worley_noise::get_rand_offset, src/noise.h:141-145 (voronoi's is the same, :173-177)
  const V rv = v3(dot(u, v3(127.1, 311.7, 74.7)), dot(u, v3(269.5, 183.3, 246.1)), dot(u, v3(113.5, 271.9, 307.7)));
  const V res = v3(sin(rv.x), sin(rv.y), sin(rv.z)) * 43758.5453;
  return res - vfloor(res);
}
DEV_COLD double worley_noise(V p, bool voronoi) { // src/noise.h:147-167, :179-200
  const V fp = vfloor(p);
  float min_dist = 3.402823466e+38f;
  float colr = 0.0f;
  for (int i = -1; i <= 1; i++)
    for (int j = -1; j <= 1; j++)
      for (int k = -1; k <= 1; k++) {
        const V cell = fp + v3(i, j, k);
        const V pos = cell + cell_offset(cell);
        const float dist = (float)length(pos - p);
        if (dist < min_dist) {
          min_dist = dist;
          if (voronoi) colr = (float)cell_offset(pos).x;
        }
      }
  return voronoi ? (double)colr : (double)(min_dist * min_dist);
}

// One texel as a colour: 8-bit images as the reference reads them (byte / 256, src/texture.h:72-73), float images (12 or 16
// bytes per pixel, the opt-in un-quantised ingest) as they are: no clamp, so an HDR environment keeps its values above 1.
// EXT (here and in tex_sample / finish_hit / shade_hit): the scene uses one of the opt-in extensions (float images, bilinear
// lookups, triangle texture coordinates).  The wavefront kernels are instantiated for both values and a scene without opt-ins
// runs the EXT = false ones, whose code is the reference path only: with the extensions compiled into the 64-register shade
// kernels, ptxas moved spills into the hot path (lambertian kernel +6.5 %, C2 -2 %; profiles/r02_ac_optins.jsonl).
template <bool EXT = true> DEV V image_texel(const DevScene &S, const b200rt_image &im, int i, int j) {
  const uint8_t *px = S.image_bytes + im.offset + ((size_t)j * im.width + i) * im.bytes_per_pixel;
  if (EXT && im.bytes_per_pixel >= 12) {
    const float *f = reinterpret_cast<const float *>(px);
    return v3((double)f[0], (double)f[1], (double)f[2]);
  }
  const double cs = 1 / 256.0;
  return v3(px[0] * cs, px[1] * cs, px[2] * cs);
}

template <bool EXT = true> DEV V tex_sample(const DevScene &S, int tex, double u, double v, V p) {
  const b200rt_texture &t = S.textures[tex];
  switch (t.kind) {
  case B200RT_TEX_SOLID: return v3(t.c0[0], t.c0[1], t.c0[2]);
  case B200RT_TEX_CHECKER: { // src/texture.h:47-56
    const V q = p / t.scale;
    const int total = (int)floor(q.x) + (int)floor(q.y) + (int)floor(q.z);
    return (total % 2 == 0) ? v3(t.c1[0], t.c1[1], t.c1[2]) : v3(t.c0[0], t.c0[1], t.c0[2]);
  }
  case B200RT_TEX_PICTURE: { // src/texture.h:67-74, image::pixel_data src/image.h:72-97
    const b200rt_image im = S.images[t.aux];
    const double cs = 1 / 256.0;
    if (!im.has_data) return v3(255 * cs, 0 * cs, 255 * cs);
    int i = (int)(im.width * u);
    int j = (int)(im.height * (1 - v));
    i = i < 0 ? 0 : (i < im.width ? i : im.width - 1);
    j = j < 0 ? 0 : (j < im.height ? j : im.height - 1);
    return image_texel<EXT>(S, im, i, j);
  }
  case B200RT_TEX_PICTURE_LINEAR: { // opt-in: bilinear between texel centres, edges clamped
    const b200rt_image im = S.images[t.aux];
    if (!EXT) return v3(0, 0, 0); // not reachable: scenes with this texture kind run the EXT kernels
    if (!im.has_data) return v3(255 / 256.0, 0.0, 255 / 256.0);
    const double x = im.width * u - 0.5, y = im.height * (1 - v) - 0.5;
    const double fx0 = floor(x), fy0 = floor(y);
    const double ax = x - fx0, ay = y - fy0;
    const int W = im.width, H = im.height;
    auto cl = [](double f, int n) { return f < 0 ? 0 : (f > n - 1 ? n - 1 : (int)f); };
    const int i0 = cl(fx0, W), i1 = cl(fx0 + 1, W), j0 = cl(fy0, H), j1 = cl(fy0 + 1, H);
    const V c00 = image_texel<EXT>(S, im, i0, j0), c10 = image_texel<EXT>(S, im, i1, j0), c01 = image_texel<EXT>(S, im, i0, j1), c11 = image_texel<EXT>(S, im, i1, j1);
    const V top = (1 - ax) * c00 + ax * c10, bot = (1 - ax) * c01 + ax * c11;
    return (1 - ay) * top + ay * bot;
  }
  case B200RT_TEX_PERLIN: { // src/texture.h:86-88
    const double *vec = S.perlin_vec + (size_t)t.aux * 768;
    const int32_t *perm = S.perlin_perm + (size_t)t.aux * 256;
    return v3(.5, .5, .5) * (1 + sin(p.x + 70 * perlin_turb(vec, perm, 7, p / t.scale)));
  }
  case B200RT_TEX_VALUE: {
    const b200rt_value_noise vn = S.value_noise[t.aux];
    const double k = value_noise(S.value_lattice + vn.offset, vn.resolution, p);
    return v3(k, k, k);
  }
  case B200RT_TEX_WORLEY: {
    const double k = worley_noise(p, false);
    return v3(k, k, k);
  }
  case B200RT_TEX_VORONOI: {
    const double k = worley_noise(p, true);
    return v3(k, k, k);
  }
  }
  return v3(0, 0, 0);
}

} // namespace b200
