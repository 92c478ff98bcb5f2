// mirror.cuh — node-for-node traversal of the reference's own scene graph (B200RT_TRAVERSAL_MIRROR).
//
// Walks the uploaded b200rt_node graph exactly as the reference walks its hittable objects: aabb::hit with
// its divisions and std::min/std::max forms (src/aabb.h:28-33, 45-69), bvh_node::hit left then right with
// the shrunk interval (src/bvh_node.h:49-59), hittable_list::hit (src/hittable_list.h:20-31), the
// translate / rotate_* wrappers (src/hittable.h:75-82, 125-149, 192-216, 259-284) and volumne::hit
// (src/volumne.h:18-46).  It reproduces even the reference's accidents (boxes that cull grazing rays,
// NaN t accepted by triangle::hit: SURVEY.md A.3, A.15) and is the arbiter for the compacted traversal.
// Device recursion: the host raises cudaLimitStackSize before launching kernels that include this header.
#pragma once
#include "intersect.cuh"

namespace b200 {

DEV double stdmin(double a, double b) { return (b < a) ? b : a; }
DEV double stdmax(double a, double b) { return (a < b) ? b : a; }

DEV bool aabb_hit(const double *f, const Ray &r, double lo, double hi) {
  double t1 = (f[0] - r.o.x) / r.d.x, t2 = (f[1] - r.o.x) / r.d.x;
  const double x0 = stdmin(t1, t2), x1 = stdmax(t1, t2);
  t1 = (f[2] - r.o.y) / r.d.y, t2 = (f[3] - r.o.y) / r.d.y;
  const double y0 = stdmin(t1, t2), y1 = stdmax(t1, t2);
  t1 = (f[4] - r.o.z) / r.d.z, t2 = (f[5] - r.o.z) / r.d.z;
  const double z0 = stdmin(t1, t2), z1 = stdmax(t1, t2);
  const double t_min = stdmax(stdmax(x0, y0), z0);
  const double t_max = stdmin(stdmin(x1, y1), z1);
  if (t_min > lo) lo = t_min;
  if (t_max < hi) hi = t_max;
  return lo < hi;
}

struct MirrorCtx {
  const DevScene *S;
  Stream st;
  int bounce;
};

__device__ __noinline__ bool mirror_hit(const MirrorCtx &cx, int idx, const Ray &r, double lo, double hi, Rec &rec, int visit);

DEV bool mirror_leaf(const MirrorCtx &cx, const b200rt_node &n, const Ray &r, double lo, double hi, Rec &rec, int visit) {
  const DevScene &S = *cx.S;
  double t, a, b;
  Entry e;
  e.kind_chain = 0;
  e.row = n.a;
  e.prim = n.c;
  e.order = 0;
  switch (n.kind) {
  case B200RT_NODE_SPHERE: {
    const bool moving = S.spheres[(size_t)n.a * B200RT_SPHERE_STRIDE + 4] != 0.0;
    if (!sphere_row_t(S, n.a, moving, r, lo, hi, t)) return false;
    e.kind_chain = moving ? E_SPHERE_MOVING : E_SPHERE;
    break;
  }
  case B200RT_NODE_QUAD:
    if (!quad_row_t(S, n.a, r, lo, hi, t, a, b)) return false;
    e.kind_chain = E_QUAD;
    break;
  case B200RT_NODE_TRIANGLE:
    if (!tri_row_t(S, n.a, r, lo, hi, t)) return false; // NaN t passes, as in the reference
    e.kind_chain = E_TRI;
    break;
  default: { // B200RT_NODE_VOLUME
    const b200rt_volume vol = S.volumes[n.a];
    Rec r1, r2;
    if (!mirror_hit(cx, vol.boundary_node, r, -B200_INF, B200_INF, r1, 0)) return false;
    if (!mirror_hit(cx, vol.boundary_node, r, r1.t + 0.0001, B200_INF, r2, 0)) return false;
    if (r1.t < lo) r1.t = lo;
    if (r2.t > hi) r2.t = hi;
    if (r1.t >= r2.t) return false;
    if (r1.t < 0) r1.t = 0;
    const double ray_length = length(r.d);
    const double inside = (r2.t - r1.t) * ray_length;
    const double rnd = cx.st.get(cx.bounce, P_VOLUME, (uint32_t)n.c * 8u + (uint32_t)visit).u0;
    const double hit_distance = -1.0 / vol.density * log(rnd);
    if (hit_distance > inside) return false;
    rec.t = r1.t + hit_distance / ray_length;
    rec.p = at(r, rec.t);
    rec.normal = v3(1, 0, 0);
    rec.front_face = true;
    rec.mat = vol.material;
    rec.u = 0;
    rec.v = 0;
    rec.prim = n.c;
    return true;
  }
  }
  finish_hit(S, e, r, t, rec); // chain 0: r is already in the leaf's frame
  return true;
}

__device__ __noinline__ bool mirror_hit(const MirrorCtx &cx, int idx, const Ray &r, double lo, double hi, Rec &rec, int visit) {
  const DevScene &S = *cx.S;
  const b200rt_node &n = S.nodes[idx];
  switch (n.kind) {
  case B200RT_NODE_BVH: {
    if (!aabb_hit(n.f, r, lo, hi)) return false;
    const bool twice = n.a == n.b;
    const bool hit_left = mirror_hit(cx, n.a, r, lo, hi, rec, twice ? visit * 2 : visit);
    const bool hit_right = mirror_hit(cx, n.b, r, lo, hit_left ? rec.t : hi, rec, twice ? visit * 2 + 1 : visit);
    return hit_left || hit_right;
  }
  case B200RT_NODE_LIST: {
    Rec tmp;
    bool any = false;
    for (int k = 0; k < n.b; k++) {
      if (mirror_hit(cx, (int)S.list_children[n.a + k], r, lo, hi, tmp, visit)) {
        any = true;
        hi = tmp.t;
        rec = tmp;
      }
    }
    return any;
  }
  case B200RT_NODE_TRANSLATE: {
    const V off = v3(n.f[0], n.f[1], n.f[2]);
    const Ray rr{r.o - off, r.d, r.tm};
    if (!mirror_hit(cx, n.a, rr, lo, hi, rec, visit)) return false;
    rec.p = rec.p + off;
    return true;
  }
  case B200RT_NODE_ROTATE_X:
  case B200RT_NODE_ROTATE_Y:
  case B200RT_NODE_ROTATE_Z: {
    Ray rr = r;
    rot_fwd(n.kind, rr.o, n.f[0], n.f[1]);
    rot_fwd(n.kind, rr.d, n.f[0], n.f[1]);
    if (!mirror_hit(cx, n.a, rr, lo, hi, rec, visit)) return false;
    rot_back(n.kind, rec.p, n.f[0], n.f[1]);
    rot_back(n.kind, rec.normal, n.f[0], n.f[1]);
    return true;
  }
  default: return mirror_leaf(cx, n, r, lo, hi, rec, visit);
  }
}

} // namespace b200
