// traverse.cuh — closest-hit search over the compacted BVH2: what replaces bvh_node::hit (src/bvh_node.h:49-59)
// and aabb::hit (src/aabb.h:28-33, 45-69) on the wavefront path.
//
// Two precisions, two jobs:
//   * BOX TESTS are a conservative binary32 filter.  One 64-byte node fetch (four 128-bit read-only loads)
//     holds both children's planes, rounded outward at build time.  Per plane one FFMA
//         t = plane * (1/d) - (o/d -/+ E)
//     where o/d is formed in binary64 and E bounds every rounding of the fp32 evaluation
//     (E = 2^-23 M / |d| per axis, M = largest |plane| of the scene), plus 4 ulp relative slack on the
//     final interval.  A box the exact ray touches is never rejected; a few untouched boxes are accepted.
//   * LEAF TESTS are binary64 in the reference's operation order (intersect.cuh), so t / p / normal and the
//     winning primitive are bit-identical to the reference whatever the filter let through.
// A candidate replaces the current best iff  t < best.t  or  (t == best.t and order > best.order): the
// reference's closed-interval "last visited wins" rule expressed without its visiting order,
// which is what makes the result independent of the tree we traverse.
#pragma once
#include "intersect.cuh"

namespace b200 {

#define B200_STACK 64
#define B200_DONE B200_EMPTY_ROOT

struct Best {
  double t;
  int32_t entry;
  uint32_t order;
};

// Per-ray constants of the fp32 box filter.  The filter works in a rescaled parameter t' = t * tsc, d' = d / tsc, with tsc the
// power of two just below the largest direction component (d' lands in [1, 2)): exact, and it keeps rays whose direction is
// uniformly tiny or huge (t of 1e30 and beyond is legitimate in the reference) inside fp32 range.
struct RaySetup {
  float ix, iy, iz; // 1/(d sc), clamped to +-2^100 (a zero component behaves like an almost-parallel ray)
  float c0x, c1x;   // what is subtracted from x0 * ix and x1 * ix: o/(d sc) with the margin on the safe side
  float c0y, c1y;
  float c0z, c1z;
  double tsc;
};

DEV void setup_axis(double o, double d, double M, float &inv_f, float &c0, float &c1) {
  const double B = 1.2676506002282294e30; // 2^100
  double inv = 1.0 / d;
  if (inv > B) inv = B;
  if (inv < -B) inv = -B;
  const double ainv = fabs(inv);
  // what the filter cannot bound (NaN anywhere, |o| beyond 2^26): plane * 0 -+ inf = this axis accepts every box
  if (!(ainv >= 0.25) || !(fabs(o) <= 6.7108864e7)) {
    inv_f = 0.0f;
    c0 = __int_as_float(0x7f800000); // plane*0 - (+inf) = -inf
    c1 = -c0;                        // plane*0 - (-inf) = +inf
    return;
  }
  const double oi = o * inv;
  const double E = 1.1920928955078125e-7 * M * ainv + 1e-37; // 2^-23 = 2 ulp32: inv_f is 1/d to half an ulp, |plane| <= M; o/d itself is rounded on the safe side below
  const float cn = __double2float_ru(oi + E); // near planes: t comes out too small, never too large
  const float cf = __double2float_rd(oi - E); // far planes: too large, never too small
  inv_f = (float)inv;
  // plane x0 is the near one when the ray travels towards +axis
  c0 = inv > 0 ? cn : cf;
  c1 = inv > 0 ? cf : cn;
}

// Largest biased exponent among the direction components, or 0 when it cannot be used as a scale (zero / denormal /
// inf / NaN directions: scale 1, and setup_axis' clamps take over).
DEV long long ray_exponent(const V &d) {
  const long long ex = (__double_as_longlong(d.x) >> 52) & 0x7ff, ey = (__double_as_longlong(d.y) >> 52) & 0x7ff, ez = (__double_as_longlong(d.z) >> 52) & 0x7ff;
  long long e = ex > ey ? ex : ey;
  e = e > ez ? e : ez;
  return (e >= 1 && e <= 2045) ? e : 0;
}
DEV double ray_tsc(const V &d) { // 2^(e - 1023)
  const long long e = ray_exponent(d);
  return e ? __longlong_as_double(e << 52) : 1.0;
}

DEV RaySetup make_setup(const double *bound, const Ray &r) {
  RaySetup s;
  const long long e = ray_exponent(r.d);
  s.tsc = e ? __longlong_as_double(e << 52) : 1.0;
  const double sc = e ? __longlong_as_double((2046LL - e) << 52) : 1.0; // 1 / tsc
  setup_axis(r.o.x, r.d.x * sc, bound[0], s.ix, s.c0x, s.c1x);
  setup_axis(r.o.y, r.d.y * sc, bound[1], s.iy, s.c0y, s.c1y);
  setup_axis(r.o.z, r.d.z * sc, bound[2], s.iz, s.c0z, s.c1z);
  return s;
}

// Conservative entry / exit parameters of one box for the interval [lo, hi] (already rounded outward to fp32): the box may be
// hit iff  tn <= tf.  tn is the plain maximum of the near planes and lo; ALL the relative slack for the FFMAs' own rounding sits
// on tf = f (1 + 1e-6).  That is enough because tn >= lo > 0 always (lo = 0.001 tsc rounded down): for f >= 0 the former
// two-sided test  n (1 - 2^-21) <= f (1 + 2^-21)  is implied by  n <= f (1 + 2^-20 + ...) <= f (1 + 1e-6), and for f < 0 both
// reject.  One FFMA per box instead of two and no |.| (r02: 2 of ~60 instructions per node step).
DEV void slab32(float x0, float x1, float y0, float y1, float z0, float z1, const RaySetup &s, float lo, float hi, float &tn, float &tf) {
  const float ax = __fmaf_rn(x0, s.ix, -s.c0x), bx = __fmaf_rn(x1, s.ix, -s.c1x);
  const float ay = __fmaf_rn(y0, s.iy, -s.c0y), by = __fmaf_rn(y1, s.iy, -s.c1y);
  const float az = __fmaf_rn(z0, s.iz, -s.c0z), bz = __fmaf_rn(z1, s.iz, -s.c1z);
  // fminf / fmaxf drop NaNs (inf * 0 on an infinite plane): such an axis does not constrain the interval
  tn = fmaxf(fmaxf(fminf(ax, bx), fminf(ay, by)), fmaxf(fminf(az, bz), lo));
  const float f = fminf(fminf(fmaxf(ax, bx), fmaxf(ay, by)), fminf(fmaxf(az, bz), hi));
  tf = __fmaf_rn(f, 1.0e-6f, f);
}

DEV void leaf_test(const DevScene &S, int32_t ref, const Ray &r, double lo, double hi, const Stream &st, int bounce, Best &best) {
  const uint32_t v = (uint32_t)~ref;
  const uint32_t first = v >> 3, count = (v & 7u) + 1u;
  for (uint32_t k = 0; k < count; k++) {
    const int4 raw = __ldg(reinterpret_cast<const int4 *>(S.entries + first + k));
    Entry e;
    e.kind_chain = (uint32_t)raw.x, e.row = raw.y, e.prim = raw.z, e.order = (uint32_t)raw.w;
    double t;
    // like the reference, later leaves see the interval shrunk to the best t so far (closed: ties get through)
    if (entry_t(S, e, r, lo, best.entry < 0 ? hi : best.t, st, bounce, t)) {
      if (best.entry < 0 || t < best.t || (t == best.t && e.order > best.order)) {
        best.t = t;
        best.entry = (int32_t)(first + k);
        best.order = e.order;
      }
    }
  }
}

// Experiment (-DB200_PUSH_PREFETCH=1): a reference that goes onto the stack will be read some node steps later, so ask for
// its line now (prefetch.global.L1: no register, no wait).  Measured on B200 (profiles/r01_o_bvh4.jsonl): SLOWER on every
// scene (Sponza stand-in 1525 vs 1597, RTNW 2004 vs 2079, RTOW 5542 vs 5617 Mrays/s): about half of the pushed references
// are culled before they are popped, and their lines evict nodes that are in use.  Off.
#ifndef B200_PUSH_PREFETCH
#define B200_PUSH_PREFETCH 0
#endif
DEV void prefetch_ref(const DevScene &S, int32_t ref, bool wide) {
#if B200_PUSH_PREFETCH == 1 && !defined(B200_HOST_EMU)
  const void *p = ref >= 0 ? (wide ? (const void *)(S.bvh4 + ref) : (const void *)(S.bvh + ref)) : (const void *)(S.entries + ((uint32_t)~ref >> 3));
  asm volatile("prefetch.global.L1 [%0];" ::"l"(p));
#endif
}

// One inner node: test both children, return the nearer accepted child (or B200_DONE) and push the other.  Written without
// control flow (selects and one predicated store) so that the node loop carries no reconvergence barriers.
DEV int32_t node_step(const DevScene &S, int32_t cur, const RaySetup &rs, float lo32, float hi32, int32_t *stack, int &sp) {
  const float4 *q = reinterpret_cast<const float4 *>(S.bvh + cur);
  const float4 a = __ldg(q), b = __ldg(q + 1), c = __ldg(q + 2);
  const int2 ch = __ldg(reinterpret_cast<const int2 *>(q + 3));
  float ln, lf, rn, rf;
  slab32(a.x, a.y, a.z, a.w, b.x, b.y, rs, lo32, hi32, ln, lf);
  slab32(b.z, b.w, c.x, c.y, c.z, c.w, rs, lo32, hi32, rn, rf);
  const bool hl = ln <= lf, hr = rn <= rf, both = hl && hr;
  const bool left_first = ln <= rn;
  const int32_t near_ref = left_first ? ch.x : ch.y, far_ref = left_first ? ch.y : ch.x;
  if (both) { // cannot overflow: b200rt_upload_scene rejects trees deeper than B200_STACK (one push per level)
    stack[sp++] = far_ref;
    prefetch_ref(S, far_ref, false);
  }
  return both ? near_ref : (hl ? ch.x : (hr ? ch.y : B200_DONE));
}

// One BVH4 node: test up to four children, return the nearest accepted one (or B200_DONE) and push the others, farthest
// first, so that they pop nearest first.
DEV void cswap4(float &ka, int32_t &ca, float &kb, int32_t &cb) {
  const bool s = kb < ka;
  const float k0 = s ? kb : ka, k1 = s ? ka : kb;
  const int32_t c0 = s ? cb : ca, c1 = s ? ca : cb;
  ka = k0, kb = k1, ca = c0, cb = c1;
}
DEV int32_t node_step4(const DevScene &S, int32_t cur, const RaySetup &rs, float lo32, float hi32, int32_t *stack, int &sp) {
  const float4 *q = reinterpret_cast<const float4 *>(S.bvh4 + cur);
  const float4 X0 = __ldg(q), X1 = __ldg(q + 1), Y0 = __ldg(q + 2), Y1 = __ldg(q + 3), Z0 = __ldg(q + 4), Z1 = __ldg(q + 5);
  const int4 ch = __ldg(reinterpret_cast<const int4 *>(q + 6));
  const float INF32 = __int_as_float(0x7f800000);
  float n0, f0, n1, f1, n2, f2, n3, f3;
  slab32(X0.x, X1.x, Y0.x, Y1.x, Z0.x, Z1.x, rs, lo32, hi32, n0, f0);
  slab32(X0.y, X1.y, Y0.y, Y1.y, Z0.y, Z1.y, rs, lo32, hi32, n1, f1);
  slab32(X0.z, X1.z, Y0.z, Y1.z, Z0.z, Z1.z, rs, lo32, hi32, n2, f2);
  slab32(X0.w, X1.w, Y0.w, Y1.w, Z0.w, Z1.w, rs, lo32, hi32, n3, f3);
  // key = entry parameter of an accepted child (kept below +inf), +inf for a rejected or absent one
  const float BIG = 3.0e38f;
  float k0 = (n0 <= f0 && ch.x != B200_DONE) ? fminf(n0, BIG) : INF32, k1 = (n1 <= f1 && ch.y != B200_DONE) ? fminf(n1, BIG) : INF32;
  float k2 = (n2 <= f2 && ch.z != B200_DONE) ? fminf(n2, BIG) : INF32, k3 = (n3 <= f3 && ch.w != B200_DONE) ? fminf(n3, BIG) : INF32;
  int32_t c0 = ch.x, c1 = ch.y, c2 = ch.z, c3 = ch.w;
  cswap4(k0, c0, k1, c1);
  cswap4(k2, c2, k3, c3);
  cswap4(k0, c0, k2, c2);
  cswap4(k1, c1, k3, c3);
  cswap4(k1, c1, k2, c2);
  // cannot overflow: the BVH4 is only used when 3 * depth <= B200_STACK (api.cu: up to three pushes per level)
#if B200_PUSH_PREFETCH == 2 // only the child that pops next (the second nearest): r02 experiment on the latency-bound Sponza stand-in
  if (k3 < INF32) stack[sp++] = c3;
  if (k2 < INF32) stack[sp++] = c2;
  if (k1 < INF32) {
    stack[sp++] = c1;
#ifndef B200_HOST_EMU
    const void *p = c1 >= 0 ? (const void *)(S.bvh4 + c1) : (const void *)(S.entries + ((uint32_t)~c1 >> 3));
    asm volatile("prefetch.global.L1 [%0];" ::"l"(p));
#endif
  }
#else
  if (k3 < INF32) stack[sp++] = c3, prefetch_ref(S, c3, true);
  if (k2 < INF32) stack[sp++] = c2, prefetch_ref(S, c2, true);
  if (k1 < INF32) stack[sp++] = c1, prefetch_ref(S, c1, true);
#endif
  return k0 < INF32 ? c0 : B200_DONE;
}

// world.hit(r, interval(lo, hi), rec) -> (entry, t); entry < 0 on a miss.
DEV Best trace_closest(const DevScene &S, const Ray &r, double lo, double hi, const Stream &st, int bounce) {
  Best best;
  best.t = hi;
  best.entry = -1;
  best.order = 0;
  if (S.bvh_root == B200_EMPTY_ROOT) return best;
  const RaySetup rs = make_setup(S.bound, r);
  const float lo32 = __double2float_rd(lo * rs.tsc);
  float hi32 = __double2float_ru(hi * rs.tsc);
  int32_t stack[B200_STACK];
  int sp = 0;
  int32_t cur = S.bvh_root;
  for (;;) {
    while (cur >= 0) { // inner nodes
      cur = S.bvh4 ? node_step4(S, cur, rs, lo32, hi32, stack, sp) : node_step(S, cur, rs, lo32, hi32, stack, sp);
      if (cur == B200_DONE && sp > 0) cur = stack[--sp];
    }
    if (cur == B200_DONE) return best;
    leaf_test(S, cur, r, lo, hi, st, bounce, best);
    if (best.entry >= 0) hi32 = __double2float_ru(best.t * rs.tsc);
    cur = sp > 0 ? stack[--sp] : B200_DONE;
  }
}

} // namespace b200
