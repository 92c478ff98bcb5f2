// traverse.cuh — closest-hit search over the compacted BVH2: what replaces bvh_node::hit (src/bvh_node.h:49-59)
// and aabb::hit (src/aabb.h:28-33, 45-69) on the wavefront path.
//
//   * one 128-byte node fetch (eight 128-bit read-only loads) tests both children;
//   * slab test on precomputed reciprocals; NaN-safe min/max (fmin/fmax drop NaNs, so a zero direction
//     component with the origin on a slab plane does not poison the interval);
//   * near child first, far child on a short per-thread stack;
//   * a candidate replaces the current best iff  t < best.t  or  (t == best.t and order > best.order):
//     the reference's closed-interval "last visited wins" rule expressed without its visiting order.
#pragma once
#include "intersect.cuh"

namespace b200 {

#define B200_STACK 64

struct Best {
  double t;
  int32_t entry;
  uint32_t order;
};

DEV void slab2(const double *bx, const V &o, const V &inv, double lo, double hi, double &tn, double &tf) {
  double x0, x1, y0, y1, z0, z1;
  ld2(bx, x0, x1);
  ld2(bx + 2, y0, y1);
  ld2(bx + 4, z0, z1);
  const double ax = (x0 - o.x) * inv.x, bx_ = (x1 - o.x) * inv.x;
  const double ay = (y0 - o.y) * inv.y, by = (y1 - o.y) * inv.y;
  const double az = (z0 - o.z) * inv.z, bz = (z1 - o.z) * inv.z;
  tn = fmax(fmax(fmin(ax, bx_), fmin(ay, by)), fmax(fmin(az, bz), lo));
  tf = fmin(fmin(fmax(ax, bx_), fmax(ay, by)), fmin(fmax(az, bz), hi));
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

// world.hit(r, interval(lo, hi), rec) -> (entry, t); entry < 0 on a miss.
DEV Best trace_closest(const DevScene &S, const Ray &r, double lo, double hi, const Stream &st, int bounce) {
  Best best;
  best.t = hi;
  best.entry = -1;
  best.order = 0;
  if (S.bvh_root == B200_EMPTY_ROOT) return best;
  if (S.bvh_root < 0) {
    leaf_test(S, S.bvh_root, r, lo, hi, st, bounce, best);
    return best;
  }
  const V inv = v3(1.0 / r.d.x, 1.0 / r.d.y, 1.0 / r.d.z);
  int32_t stack[B200_STACK];
  int sp = 0;
  int32_t cur = S.bvh_root;
  for (;;) {
    const BvhNode *nd = S.bvh + cur;
    double ln, lf, rn, rf;
    const double far_limit = best.entry < 0 ? hi : best.t;
    slab2(nd->lbox, r.o, inv, lo, far_limit, ln, lf);
    slab2(nd->rbox, r.o, inv, lo, far_limit, rn, rf);
    const int2 ch = __ldg(reinterpret_cast<const int2 *>(&nd->left));
    const bool hl = ln <= lf, hr = rn <= rf;
    int32_t next = B200_EMPTY_ROOT;
    if (hl && hr) {
      const bool left_first = ln <= rn;
      next = left_first ? ch.x : ch.y;
      const int32_t other = left_first ? ch.y : ch.x;
      if (sp < B200_STACK) stack[sp++] = other;
    } else if (hl) {
      next = ch.x;
    } else if (hr) {
      next = ch.y;
    }
    // descend; leaves are tested as they are reached
    for (;;) {
      if (next == B200_EMPTY_ROOT) {
        if (sp == 0) return best;
        next = stack[--sp];
      }
      if (next >= 0) break;
      leaf_test(S, next, r, lo, hi, st, bounce, best);
      next = B200_EMPTY_ROOT;
    }
    cur = next;
  }
}

} // namespace b200
