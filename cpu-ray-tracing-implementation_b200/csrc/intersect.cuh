// intersect.cuh — primitive intersection in binary64, reference operation order.
//
//   sphere     src/sphere.h:40-74      quad      src/quad.h:30-64
//   triangle   src/triangle.h:8-40     volumne   src/volumne.h:18-46
//   translate / rotate_x / rotate_y / rotate_z   src/hittable.h:75-82, 125-149, 192-216, 259-284
//
// The wavefront splits hittable::hit in two: `entry_t` answers "at which t does this leaf accept the ray"
// (all extend needs while it searches for the closest hit), `finish_hit` rebuilds the whole hit_record
// (p, normal, u, v, front_face, material) from (leaf, t) when shading needs it.  Both run the
// reference's operations in the reference's order, so the record is the one bvh_node::hit would return.
#pragma once
#include "scene_dev.cuh"

namespace b200 {

DEV bool contains(double lo, double hi, double x) { return lo <= x && x <= hi; } // src/interval.h:14

// ---------------------------------------------------------------- transform chains
// rotate_x / rotate_y / rotate_z act on the coordinate pairs (y,z) / (x,z) / (x,y) (src/hittable.h:125-149, 192-216, 259-284).
// Written per axis with named components: indexing a V by a run-time axis number expands into select chains that made
// the transform code 10 000 of the extend kernel's 23 000 instructions (nvdisasm line table, profiles/r02_m_codesize.txt).
DEV void rot_fwd(int kind, V &a, double s, double c) {
  if (kind == B200RT_NODE_ROTATE_X) {
    const double p0 = a.y, p1 = a.z;
    a.y = c * p0 - s * p1;
    a.z = s * p0 + c * p1;
  } else if (kind == B200RT_NODE_ROTATE_Y) {
    const double p0 = a.x, p1 = a.z;
    a.x = c * p0 - s * p1;
    a.z = s * p0 + c * p1;
  } else {
    const double p0 = a.x, p1 = a.y;
    a.x = c * p0 - s * p1;
    a.y = s * p0 + c * p1;
  }
}
DEV void rot_back(int kind, V &a, double s, double c) {
  if (kind == B200RT_NODE_ROTATE_X) {
    const double p0 = a.y, p1 = a.z;
    a.y = c * p0 + s * p1;
    a.z = -s * p0 + c * p1;
  } else if (kind == B200RT_NODE_ROTATE_Y) {
    const double p0 = a.x, p1 = a.z;
    a.x = c * p0 + s * p1;
    a.z = -s * p0 + c * p1;
  } else {
    const double p0 = a.x, p1 = a.y;
    a.x = c * p0 + s * p1;
    a.y = -s * p0 + c * p1;
  }
}
// world ray -> object ray, outermost transform first
DEV Ray chain_apply(const DevScene &S, int chain, Ray r) {
  if (chain == 0) return r;
  const Chain ch = S.chains[chain];
#pragma unroll 1
  for (int k = 0; k < ch.count; k++) {
    const ChainOp op = S.chain_ops[ch.first + k];
    if (op.kind == B200RT_NODE_TRANSLATE) {
      r.o = r.o - v3(op.f0, op.f1, op.f2);
    } else {
      rot_fwd(op.kind, r.o, op.f0, op.f1);
      rot_fwd(op.kind, r.d, op.f0, op.f1);
    }
  }
  return r;
}
// object-space p / normal -> world, innermost transform first
DEV void chain_unapply(const DevScene &S, int chain, V &p, V &n) {
  if (chain == 0) return;
  const Chain ch = S.chains[chain];
#pragma unroll 1
  for (int k = ch.count - 1; k >= 0; k--) {
    const ChainOp op = S.chain_ops[ch.first + k];
    if (op.kind == B200RT_NODE_TRANSLATE) {
      p = p + v3(op.f0, op.f1, op.f2);
    } else {
      rot_back(op.kind, p, op.f0, op.f1);
      rot_back(op.kind, n, op.f0, op.f1);
    }
  }
}

// ---------------------------------------------------------------- 128-bit row loads
DEV void ld2(const double *p, double &a, double &b) {
  const double2 q = __ldg(reinterpret_cast<const double2 *>(p));
  a = q.x;
  b = q.y;
}

// ---------------------------------------------------------------- sphere
// The root that sphere::hit returns for [lo, hi]: the near one if inside, else the far one.
DEV bool sphere_t(V center, double radius, const Ray &r, double lo, double hi, double &t) {
  const V oc = r.o - center;
  const double a = dot(r.d, r.d);
  const double b = 2.0 * dot(r.d, oc);
  const double c = dot(oc, oc) - radius * radius;
  const double disc = b * b - 4 * a * c;
  if (disc < 0) return false;
  const double sq = sqrt(disc);
  const double den = 2.0 * a;
  // root = num / den with den = 2 (d.d) > 0: a numerator beyond bound * den by more than 1e-9 relative is a certain
  // "outside [lo, hi]" without dividing (same argument as in tri_row_t); everything else takes the reference's division.
  const double K = 1e-9, l = lo * den, h = hi * den;
  const bool usable = den > 1e-200 && den < 1e100;
  double num = -b - sq;
  bool out = usable && fabs(num) > 1e-200 && (num < l - K * fabs(l) || num > h + K * fabs(h));
  double root = 0;
  if (!out) {
    root = num / den;
    out = !contains(lo, hi, root);
  }
  if (out) {
    num = -b + sq;
    if (usable && fabs(num) > 1e-200 && (num < l - K * fabs(l) || num > h + K * fabs(h))) return false;
    root = num / den;
    if (!contains(lo, hi, root)) return false;
  }
  t = root;
  return true;
}
DEV bool sphere_row_t(const DevScene &S, int row, bool moving, const Ray &r, double lo, double hi, double &t) {
  const double *q = S.spheres + (size_t)row * B200RT_SPHERE_STRIDE;
  double cx, cy, cz, rad;
  ld2(q, cx, cy);
  ld2(q + 2, cz, rad);
  V center = v3(cx, cy, cz);
  if (moving) { // src/sphere.h:83: center(time) = begin + time * (end - begin)
    const V c0 = v3(q[5], q[6], q[7]), c1 = v3(q[8], q[9], q[10]);
    center = c0 + r.tm * (c1 - c0);
  }
  return sphere_t(center, rad, r, lo, hi, t);
}
DEV void get_sphere_uv(V p, double &u, double &v) { // src/sphere.h:90-95
  const double theta = acos(-p.y);
  const double phi = atan2(-p.z, p.x) + B200_PI;
  u = phi / (2 * B200_PI);
  v = theta / B200_PI;
}
DEV void set_face_normal(Rec &rec, const Ray &r, V outward) { // src/hittable.h:26-29
  rec.front_face = dot(r.d, outward) < 0.0;
  rec.normal = rec.front_face ? outward : -outward;
}

// ---------------------------------------------------------------- quad
DEV bool quad_row_t(const DevScene &S, int row, const Ray &r, double lo, double hi, double &t, double &alpha, double &beta) {
  const double *q = S.quads + (size_t)row * B200RT_QUAD_STRIDE;
  double c0, c1, c2, u0, u1, u2, v0, v1, v2, n0, n1, n2, w0, w1, w2, D;
  ld2(q + 8, v2, n0);
  ld2(q + 10, n1, n2);
  ld2(q + 14, w2, D);
  const V unorm = v3(n0, n1, n2);
  const double qn = D - dot(unorm, r.o), qd = dot(unorm, r.d);
  { // certain "t outside [lo, hi]" from numerator and denominator, without the division (see tri_row_t)
    const double K = 1e-9, ad = fabs(qd), sn = qd < 0 ? -qn : qn;
    if (ad > 1e-200 && ad < 1e100 && fabs(sn) > 1e-200) {
      const double l = lo * ad, h = hi * ad;
      if (sn < l - K * fabs(l) || sn > h + K * fabs(h)) return false;
    }
  }
  const double tt = qn / qd;
  if (!contains(lo, hi, tt)) return false;
  ld2(q, c0, c1);
  ld2(q + 2, c2, u0);
  ld2(q + 4, u1, u2);
  ld2(q + 6, v0, v1);
  ld2(q + 12, w0, w1);
  const V p = at(r, tt) - v3(c0, c1, c2);
  const V w = v3(w0, w1, w2);
  const double a = dot(w, cross(p, v3(v0, v1, v2)));
  const double b = dot(w, cross(v3(u0, u1, u2), p));
  if (!contains(0.0, 1.0, a) || !contains(0.0, 1.0, b)) return false; // is_interior, src/quad.h:58-64
  t = tt;
  alpha = a;
  beta = b;
  return true;
}

// ---------------------------------------------------------------- triangle
DEV bool tri_core_t(const V &p0, const V &p1, const V &p2, const Ray &r, double lo, double hi, double &t);
DEV void tri32_unpack(const float4 &a, const float4 &b, const float4 &c, V &p0, V &p1, V &p2) {
  p0 = v3((double)a.x, (double)a.y, (double)a.z), p1 = v3((double)a.w, (double)b.x, (double)b.y), p2 = v3((double)b.z, (double)b.w, (double)c.x);
}
DEV bool tri_row_t(const DevScene &S, int row, const Ray &r, double lo, double hi, double &t) {
  V p0, p1, p2;
  if (S.tri32) { // 48-byte binary32 row, widened: the same binary64 operands as the 96-byte row holds (checked at upload)
    const float4 *q = S.tri32 + (size_t)row * 3;
    const float4 a = __ldg(q), b = __ldg(q + 1), c = __ldg(q + 2);
    tri32_unpack(a, b, c, p0, p1, p2);
  } else {
    const double *q = S.tris + (size_t)row * B200RT_TRI_STRIDE;
    double a0, a1, a2, b0_, b1_, b2_, c0, c1, c2, nx;
    ld2(q, a0, a1);
    ld2(q + 2, a2, b0_);
    ld2(q + 4, b1_, b2_);
    ld2(q + 6, c0, c1);
    ld2(q + 8, c2, nx);
    p0 = v3(a0, a1, a2), p1 = v3(b0_, b1_, b2_), p2 = v3(c0, c1, c2);
  }
  return tri_core_t(p0, p1, p2, r, lo, hi, t);
}
DEV bool tri_core_t(const V &p0, const V &p1, const V &p2, const Ray &r, double lo, double hi, double &t) {
  const V e1 = p1 - p0, e2 = p2 - p0, sv = r.o - p0;
  const V s1 = cross(r.d, e2), s2 = cross(sv, e1);
  const double det = dot(s1, e1);
  const double nt = dot(s2, e2), n0 = dot(s1, sv), n1 = dot(s2, r.d);
  // The reference divides all three numerators by det and then tests (src/triangle.h:8-15, 32-34).  Most candidates fail,
  // and a CERTAIN failure can be read off the numerators: the quotient's sign is exact, and a numerator beyond its bound
  // times det by more than 1e-9 relative (seven orders above the rounding of the products and of the division) is beyond it
  // after the division too.  NaN / zero det make every comparison false and fall through to the reference's own arithmetic.
  {
    const double K = 1e-9, tiny = 1e-200, ad = fabs(det);
    if (ad > tiny && ad < 1e100) { // quotients of numerators above `tiny` cannot underflow to a signed zero
      const double st = det < 0 ? -nt : nt, s0 = det < 0 ? -n0 : n0, s1n = det < 0 ? -n1 : n1; // det's sign folded in
      if (s0 < -tiny || s1n < -tiny) return false;      // b0 < 0 or b1 < 0
      if (s0 + s1n > ad + K * ad) return false;          // b0 + b1 > 1 (both terms are >= -tiny here: no cancellation)
      if (fabs(st) > tiny) {
        const double l = lo * ad, h = hi * ad;           // +-inf bounds stay +-inf and never reject
        if (st < l - K * fabs(l)) return false;          // t < lo
        if (st > h + K * fabs(h)) return false;          // t > hi
      }
    }
  }
  const double tt = nt / det, b0 = n0 / det, b1 = n1 / det;
  if (tt < lo || tt > hi) return false;
  if (b0 < 0 || b1 < 0 || b0 + b1 > 1) return false;
  t = tt;
  return true;
}

// ---------------------------------------------------------------- volume boundary: closest t over its leaves
DEV bool boundary_t(const DevScene &S, const VolInst &vi, const Ray &r, double lo, double hi, double &tout) {
  bool any = false;
  double best = hi;
  for (int k = 0; k < vi.bcount; k++) {
    const Entry e = S.bentries[vi.bfirst + k];
    const Ray ro = chain_apply(S, entry_chain(e), r);
    double t, a, b;
    bool ok;
    switch (entry_kind(e)) {
    case E_SPHERE: ok = sphere_row_t(S, e.row, false, ro, lo, best, t); break;
    case E_SPHERE_MOVING: ok = sphere_row_t(S, e.row, true, ro, lo, best, t); break;
    case E_QUAD: ok = quad_row_t(S, e.row, ro, lo, best, t, a, b); break;
    case E_TRI: ok = tri_row_t(S, e.row, ro, lo, best, t) && t == t; break;
    default: ok = false; break;
    }
    if (ok) {
      any = true;
      best = t;
    }
  }
  tout = best;
  return any;
}

// volumne::hit, src/volumne.h:18-46.  `r` is the ray in the volume's own frame.
// DEV_COLD = inlined by default (see vec.cuh: the out-of-line variant was measured slower); takes the scene through
// DevScene::self so that the experiment does not force a local-memory copy of the kernel parameter.
DEV_COLD bool volume_t(const DevScene *Sp, const Entry &e, const Ray &r, double lo, double hi, const Stream &st, int bounce, double &t) {
  const DevScene &S = *Sp;
  const VolInst vi = S.vol_inst[e.row];
  // the two boundary queries of src/volumne.h:21-24 (universe, then t1 + 0.0001 .. infinity) as ONE call site in a loop:
  // boundary_t carries every primitive test, and the extend kernel is sensitive to its instruction footprint
  double tb[2], from = -B200_INF;
#pragma unroll 1
  for (int k = 0; k < 2; k++) {
    if (!boundary_t(S, vi, r, from, B200_INF, tb[k])) return false;
    from = tb[0] + 0.0001;
  }
  double t1 = tb[0], t2 = tb[1];
  if (t1 < lo) t1 = lo;
  if (t2 > hi) t2 = hi;
  if (t1 >= t2) return false;
  if (t1 < 0) t1 = 0;
  const double ray_length = length(r.d);
  const double inside = (t2 - t1) * ray_length;
  const double rnd = st.get(bounce, P_VOLUME, (uint32_t)e.prim * 8u + (uint32_t)vi.visit).u0;
  const double hit_distance = -1.0 / S.volumes[vi.volume].density * log(rnd);
  if (hit_distance > inside) return false;
  t = t1 + hit_distance / ray_length;
  return true;
}

#ifndef B200_SPHERE_FASTPATH
#define B200_SPHERE_FASTPATH 0
#endif
// ---------------------------------------------------------------- one leaf entry: accepted t in [lo, hi]
DEV bool entry_t(const DevScene &S, const Entry &e, const Ray &rw, double lo, double hi, const Stream &st, int bounce, double &t) {
  const int kind = entry_kind(e);
#if B200_SPHERE_FASTPATH
  if (e.kind_chain == (uint32_t)E_SPHERE) return sphere_row_t(S, e.row, false, rw, lo, hi, t); // hot path: untransformed static sphere
#endif
  const Ray r = chain_apply(S, entry_chain(e), rw); // chain 0 (no transform) returns at once
  double a, b;
  switch (kind) {
  case E_SPHERE:
  case E_SPHERE_MOVING: return sphere_row_t(S, e.row, kind == E_SPHERE_MOVING, r, lo, hi, t); // one inlined copy for both
  case E_QUAD: return quad_row_t(S, e.row, r, lo, hi, t, a, b);
  case E_TRI: return tri_row_t(S, e.row, r, lo, hi, t) && t == t; // a NaN t is a "hit" in the reference; only the mirror traversal reproduces that
  case E_VOLUME: return volume_t(S.self, e, r, lo, hi, st, bounce, t);
  }
  return false;
}

// ---------------------------------------------------------------- (leaf, t) -> hit_record
// uv_mode: 1 = always fill rec.u, rec.v (b200rt_trace); 2 = only when the material's texture reads them (picture_texture is
// the only texture::sample that looks at u, v: src/texture.h:12-119), which spares the shading kernels acos + atan2 per sphere hit.
#define B200_UV_ALWAYS 1
#define B200_UV_IF_USED 2
template <bool EXT = true> DEV bool material_reads_uv(const DevScene &S, int mat) {
  const int k = S.textures[S.materials[mat].tex].kind;
  return k == B200RT_TEX_PICTURE || (EXT && k == B200RT_TEX_PICTURE_LINEAR);
}

template <bool EXT = true> DEV void finish_hit(const DevScene &S, const Entry &e, const Ray &rw, double t, Rec &rec, int uv_mode = B200_UV_ALWAYS) {
  const int chain = entry_chain(e);
  const Ray r = chain_apply(S, chain, rw);
  rec.t = t;
  rec.prim = e.prim;
  rec.p = at(r, t);
  rec.u = 0; // triangle::hit and volumne::hit do not write u, v (src/triangle.h:35-39); the ABI defines 0
  rec.v = 0;
  switch (entry_kind(e)) {
  case E_SPHERE:
  case E_SPHERE_MOVING: {
    const double *q = S.spheres + (size_t)e.row * B200RT_SPHERE_STRIDE;
    const V outward = (rec.p - v3(q[0], q[1], q[2])) / q[3]; // center_, also for moving spheres: src/sphere.h:69
    rec.mat = S.sphere_mat[e.row];
    if (uv_mode == B200_UV_ALWAYS || material_reads_uv<EXT>(S, rec.mat)) get_sphere_uv(outward, rec.u, rec.v);
    set_face_normal(rec, r, outward);
    break;
  }
  case E_QUAD: {
    const double *q = S.quads + (size_t)e.row * B200RT_QUAD_STRIDE;
    const V p = rec.p - v3(q[0], q[1], q[2]);
    const V w = v3(q[12], q[13], q[14]);
    rec.mat = S.quad_mat[e.row];
    if (uv_mode == B200_UV_ALWAYS || material_reads_uv<EXT>(S, rec.mat)) {
      rec.u = dot(w, cross(p, v3(q[6], q[7], q[8])));
      rec.v = dot(w, cross(v3(q[3], q[4], q[5]), p));
    }
    set_face_normal(rec, r, v3(q[9], q[10], q[11]));
    break;
  }
  case E_TRI: {
    const double *q = S.tris + (size_t)e.row * B200RT_TRI_STRIDE;
    set_face_normal(rec, r, v3(q[9], q[10], q[11]));
    rec.mat = S.tri_mat[e.row];
    if (EXT && S.tri_uv && (uv_mode == B200_UV_ALWAYS || material_reads_uv<EXT>(S, rec.mat))) {
      // opt-in: vertex texture coordinates weighted with moller_trumbore's b0, b1 (src/triangle.h:8-15), recomputed for the hit
      const V p0 = v3(q[0], q[1], q[2]), p1 = v3(q[3], q[4], q[5]), p2 = v3(q[6], q[7], q[8]);
      const V e1 = p1 - p0, e2 = p2 - p0, sv = r.o - p0;
      const V s1 = cross(r.d, e2), s2 = cross(sv, e1);
      const double det = dot(s1, e1), b0 = dot(s1, sv) / det, b1 = dot(s2, r.d) / det;
      const double *w = S.tri_uv + (size_t)e.row * 6;
      rec.u = (1 - b0 - b1) * w[0] + b0 * w[2] + b1 * w[4];
      rec.v = (1 - b0 - b1) * w[1] + b0 * w[3] + b1 * w[5];
    }
    break;
  }
  default: { // E_VOLUME, src/volumne.h:41-44
    const VolInst vi = S.vol_inst[e.row];
    rec.normal = v3(1, 0, 0);
    rec.front_face = true;
    rec.mat = S.volumes[vi.volume].material;
    break;
  }
  }
  chain_unapply(S, chain, rec.p, rec.normal);
}

} // namespace b200
