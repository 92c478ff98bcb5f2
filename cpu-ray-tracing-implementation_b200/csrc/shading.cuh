// shading.cuh — one level of camera::ray_color (src/camera.h:193-241) as a path-state update.
//
//   material::scatter / p_scattered / emitted   src/material.h:57-219
//   pdf::value / generate, dual_pdf, hittable_pdf   src/pdf.h:15-61, src/hittable_list.h:39-50
//   quad / sphere pdf_value / random   src/quad.h:66-78, src/sphere.h:76-81
//   camera::miss   src/camera.h:180-190
//
// The reference recurses: L = Le + atten * p_scattered * L_in / pdf.  Only path ends (a light, the
// background, depth exhaustion) produce radiance, so the recursion unrolls into a running
// throughput T and one accumulation of T * L_end per sample.
#pragma once
#include "intersect.cuh"
#include "texture.cuh"

namespace b200 {

DEV V random_in_unit_sphere(double u1, double u2) { // src/utility.h:30-42: a point ON the sphere, pole on y
  const double cos_theta = 1 - 2 * u1;
  const double sin_theta = sqrt(1 - cos_theta * cos_theta);
  const double phi = 2 * B200_PI * u2;
  double s, c;
  sincos(phi, &s, &c);
  return v3(sin_theta * c, cos_theta, sin_theta * s);
}
DEV V random_cosine_direction(double r1, double r2) { // src/utility.h:61-69, lobe about +y
  const double phi = 2 * B200_PI * r1;
  double s, c;
  sincos(phi, &s, &c);
  return v3(c * sqrt(r2), sqrt(1 - r2), s * sqrt(r2));
}
struct Onb { // src/onb.h:18-28, y = normal
  V x, y, z;
  DEV explicit Onb(V normal) {
    y = unit_vector(normal);
    const V a = (fabs(y.x) > 0.9) ? v3(0, 0, 1) : v3(1, 0, 0);
    z = unit_vector(cross(y, a));
    x = cross(y, z);
  }
  DEV V transform(V v) const { return v3(0, 0, 0) + v.x * x + v.y * y + v.z * z; } // src/onb.h:6
};
DEV V reflect(V v, V n) { return v - 2 * dot(v, n) * n; } // src/utility.h:70
DEV V refract(V v, V n, double eta) {                    // src/utility.h:71-76
  const double cos_theta = fmin(dot(-v, n), 1.0);
  const V perp = eta * (v + cos_theta * n);
  const V par = -sqrt(fabs(1.0 - length_squared(perp))) * n;
  return perp + par;
}
DEV double cosine_pdf_value(const Onb &uvw, V direction) { // src/pdf.h:37-40
  const double c = dot(unit_vector(direction), uvw.y);
  return fmax(0.0, c / B200_PI);
}

// hittable_pdf over the `light` argument of camera::render
DEV double light_pdf_value(const DevScene &S, V origin, V direction) {
  switch (S.light_kind) {
  case B200RT_LIGHT_QUAD: { // src/quad.h:66-73: one more quad::hit, no occlusion test
    const Ray r{origin, direction, 0.0};
    double t, a, b;
    if (!quad_row_t(S, S.light_index, r, 0.001, B200_INF, t, a, b)) return 0;
    const double *q = S.quads + (size_t)S.light_index * B200RT_QUAD_STRIDE;
    const V unorm = v3(q[9], q[10], q[11]);
    const V n = dot(direction, unorm) < 0.0 ? unorm : -unorm; // rec.normal after set_face_normal
    const double distance_squared = t * t * length_squared(direction);
    const double cosine = fabs(dot(unit_vector(direction), n));
    return distance_squared / (cosine * q[16]);
  }
  case B200RT_LIGHT_SPHERE: { // src/sphere.h:76-78
    const double *q = S.spheres + (size_t)S.light_index * B200RT_SPHERE_STRIDE;
    return q[3] * q[3] * B200_PI / length_squared(origin - v3(q[0], q[1], q[2]));
  }
  default: return 0.0; // src/hittable.h:39
  }
}
DEV V light_random(const DevScene &S, V origin, double u1, double u2) {
  switch (S.light_kind) {
  case B200RT_LIGHT_QUAD: { // src/quad.h:75-78
    const double *q = S.quads + (size_t)S.light_index * B200RT_QUAD_STRIDE;
    const V p = v3(q[0], q[1], q[2]) + (u1 * v3(q[3], q[4], q[5])) + (u2 * v3(q[6], q[7], q[8]));
    return p - origin;
  }
  case B200RT_LIGHT_SPHERE: { // src/sphere.h:81
    const double *q = S.spheres + (size_t)S.light_index * B200RT_SPHERE_STRIDE;
    return random_in_unit_sphere(u1, u2) * q[3];
  }
  default: return v3(1, 0, 0); // src/hittable.h:41
  }
}

// camera::miss: a unit sphere around the ray origin gives (u, v, p) for the background texture
template <bool EXT = true> DEV V miss_color(const DevScene &S, const Ray &r) {
  if (S.background_tex < 0) return v3(0, 0, 0);
  double t;
  if (!sphere_t(r.o, 1.0, r, 0.001, B200_INF, t)) return v3(0, 0, 0);
  if (S.textures[S.background_tex].kind == B200RT_TEX_SOLID) {
    const b200rt_texture &bt = S.textures[S.background_tex];
    return v3(bt.c0[0], bt.c0[1], bt.c0[2]);
  }
  const V p = at(r, t);
  const V outward = (p - r.o) / 1.0;
  double u, v;
  get_sphere_uv(outward, u, v);
  return tex_sample<EXT>(S, S.background_tex, u, v, p);
}

// Outcome of shading one hit.
struct Scatter {
  bool ends;   // path ends here; `value` is what it contributes (times the incoming throughput)
  V value;     // ends: emitted radiance; else: throughput factor
  V dir;       // direction of the scattered ray (origin = rec.p, time = incoming time)
};

// The factors ray_color combines (src/camera.h:202-240), kept apart for the parity entry b200rt_scatter.
struct ScatterDetail {
  int mode;       // 0 = scatter() returned false (emission only), 1 = kDetermined, 2 = kRandom
  V emitted;      // material::emitted
  V attenuation;  // scatter_record.attenuation
  double pdf_value, p_scattered;
};
// Draws handed in by the caller instead of Philox (b200rt_scatter): m = what material::scatter consumes, s = what the pdf does.
struct FixedDraws {
  Draws m, s;
  DEV Draws get(int, int purpose, uint32_t) const { return purpose == P_MATERIAL ? m : s; }
};

// One level of ray_color below the hit test (src/camera.h:202-240).  `b` = scatters so far on this path.
// KIND >= 0: the caller knows the material kind at compile time (material-sorted queues), so only that branch is compiled in.
// STREAM: where the draws come from (Stream = Philox by path address).  det: optional breakdown (constant nullptr on the hot path).
template <int KIND = -1, class STREAM = Stream, bool EXT = true>
DEV Scatter shade_hit(const DevScene &S, const Rec &rec, const Ray &r, const STREAM &st, int b, ScatterDetail *det = nullptr) {
  Scatter out;
  if (det) det->mode = 0, det->emitted = v3(0, 0, 0), det->attenuation = v3(0, 0, 0), det->pdf_value = 0, det->p_scattered = 0;
  b200rt_material m = S.materials[rec.mat];
  if (KIND >= 0) m.kind = KIND;
  if (m.kind == B200RT_MAT_DIFFUSE_LIGHT) { // src/material.h:211-215; scatter() is the base's false (:43)
    out.ends = true;
    out.value = rec.front_face ? tex_sample<EXT>(S, m.tex, rec.u, rec.v, rec.p) : v3(0, 0, 0);
    out.dir = v3(0, 0, 0);
    if (det) det->emitted = out.value;
    return out;
  }
  out.ends = false;
  bool determined = false;
  V attenuation = v3(0, 0, 0);
  int pdf_kind = 0; // 1 cosine about the normal, 2 uniform sphere
  switch (m.kind) {
  case B200RT_MAT_LAMBERTIAN: // src/material.h:62-67
    attenuation = tex_sample<EXT>(S, m.tex, rec.u, rec.v, rec.p);
    pdf_kind = 1;
    break;
  case B200RT_MAT_METAL: { // src/material.h:85-92: always scatters, fuzz draw taken even when fuzz == 0
    const Draws dm = st.get(b, P_MATERIAL, 0);
    V dir = unit_vector(reflect(r.d, rec.normal));
    dir = dir + (double)m.p0 * unit_vector(random_in_unit_sphere(dm.u1, dm.u2));
    determined = true;
    out.dir = dir;
    attenuation = tex_sample<EXT>(S, m.tex, rec.u, rec.v, rec.p);
    break;
  }
  case B200RT_MAT_DIELECTRIC: { // src/material.h:113-131, Schlick :135-139
    determined = true;
    attenuation = tex_sample<EXT>(S, m.tex, rec.u, rec.v, rec.p);
    const double ri = rec.front_face ? (1.0 / m.p0) : m.p0;
    const V ud = unit_vector(r.d);
    const double cos_theta = fmin(dot(-ud, rec.normal), 1.0);
    const double sin_theta = sqrt(1.0 - cos_theta * cos_theta);
    bool do_reflect = ri * sin_theta > 1.0;
    if (!do_reflect) {
      double r0 = (1 - ri) / (1 + ri);
      r0 = r0 * r0;
      const double refl = r0 + (1 - r0) * pow((1 - cos_theta), 5.0);
      do_reflect = refl > st.get(b, P_MATERIAL, 0).u0;
    }
    out.dir = do_reflect ? reflect(ud, rec.normal) : refract(ud, rec.normal, ri);
    break;
  }
  case B200RT_MAT_GLOSS: { // src/material.h:158-174
    const Draws dm = st.get(b, P_MATERIAL, 0);
    if (dm.u0 <= (double)m.p1) {
      const V specular_dir = reflect(r.d, rec.normal);
      const V diffuse_dir = Onb(rec.normal).transform(random_cosine_direction(dm.u1, dm.u2));
      const double t = (double)m.p0;
      out.dir = unit_vector((1 - t) * diffuse_dir + t * specular_dir);
      determined = true;
      attenuation = v3(1.0, 1.0, 1.0);
    } else {
      attenuation = tex_sample<EXT>(S, m.tex, rec.u, rec.v, rec.p);
      pdf_kind = 1;
    }
    break;
  }
  default: // B200RT_MAT_ISOTROPIC, src/material.h:193-200
    attenuation = tex_sample<EXT>(S, m.tex, rec.u, rec.v, rec.p);
    pdf_kind = 2;
    break;
  }
  if (det) det->attenuation = attenuation;
  if (determined) { // src/camera.h:210-214
    out.value = attenuation;
    if (det) det->mode = 1;
    return out;
  }
  // kRandom, src/camera.h:217-240
  const Draws ds = st.get(b, P_SAMPLE, 0);
  const Onb uvw(pdf_kind == 1 ? rec.normal : v3(0, 1, 0));
  V dir;
  const bool from_light = S.light_kind != B200RT_LIGHT_NONE && ds.u0 < 0.5; // dual_pdf::generate, src/pdf.h:52-57
  if (from_light)
    dir = light_random(S, rec.p, ds.u1, ds.u2);
  else if (pdf_kind == 1)
    dir = uvw.transform(random_cosine_direction(ds.u1, ds.u2)); // src/pdf.h:41
  else
    dir = unit_vector(random_in_unit_sphere(ds.u1, ds.u2)); // src/pdf.h:19
  const double mat_value = pdf_kind == 1 ? cosine_pdf_value(uvw, dir) : 1 / (4 * B200_PI);
  double pdf_value = mat_value;
  if (S.light_kind != B200RT_LIGHT_NONE) pdf_value = 0.5 * light_pdf_value(S, rec.p, dir) + 0.5 * mat_value;
  double p_scattered;
  if (pdf_kind == 1) { // src/material.h:69-72, 176-179: normal not renormalised
    const double cosine = dot(rec.normal, unit_vector(dir));
    p_scattered = cosine < 0 ? 0 : cosine / B200_PI;
  } else {
    p_scattered = 1 / (4 * B200_PI);
  }
  out.dir = dir;
  out.value = (attenuation * p_scattered) / pdf_value;
  if (det) det->mode = 2, det->pdf_value = pdf_value, det->p_scattered = p_scattered;
  return out;
}

} // namespace b200
