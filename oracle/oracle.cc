// oracle.cc — TEST INFRASTRUCTURE: a CPU restatement of the reference's per-pixel hot path
// (camera::render sample loop + recursive ray_color and everything under them) operating on the
// flattened description of include/b200rt.h.  Never linked into, imported by or executed from the
// product path; only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may load it.
//
// Parity status: PINNED.  The reference ships no golden vectors, so this file is
// pinned against the reference itself compiled here (oracle/_ref, built from /root/reference/src):
// tests/test_oracle_golden.py requires bit-identical hit records (id, t, u, v, p, normal,
// front_face), camera rays and texture lookups, and statistically identical images; the committed
// fixtures under tests/golden/ carry the reference's outputs to machines without /root/reference.
//
// Every function cites the reference lines it restates.  Arithmetic is binary64 in the reference's
// operation order (build: -O2 -ffp-contract=off, no -march => no FMA contraction).
//
// The one deliberate difference is the random source: the reference draws from the global
// std::rand stream (src/utility.h:20); here every draw is a Philox4x32-10 output addressed by
// (seed; pixel, sample, bounce, purpose, aux), the same addressing the CUDA path uses, so that the
// two can be compared sample by sample.  Distributions are the reference's.
#include <algorithm>
#include <atomic>
#include <chrono>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <limits>
#include <thread>
#include <vector>

#include "b200rt.h"

namespace {

const double infinity = std::numeric_limits<double>::infinity(); // src/utility.h:14
const double pi = 3.1415926535897932385;                         // src/utility.h:15

// ------------------------------------------------------------------ vec3 (src/vec3.h)
struct V {
  double x, y, z;
};
inline V v3(double a, double b, double c) { return {a, b, c}; }
inline V operator+(V a, V b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }  // src/vec3.h:63
inline V operator-(V a, V b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }  // src/vec3.h:65
inline V operator*(V a, V b) { return {a.x * b.x, a.y * b.y, a.z * b.z}; }  // src/vec3.h:67
inline V operator*(V a, double c) { return {a.x * c, a.y * c, a.z * c}; }   // src/vec3.h:69
inline V operator*(double c, V a) { return {a.x * c, a.y * c, a.z * c}; }   // src/vec3.h:71
inline V operator/(V a, double c) { return {a.x / c, a.y / c, a.z / c}; }   // src/vec3.h:75
inline V operator-(V a) { return {-a.x, -a.y, -a.z}; }                      // src/vec3.h:15
inline double dot(V a, V b) { return a.x * b.x + a.y * b.y + a.z * b.z; }   // src/vec3.h:61
inline V cross(V a, V b) { return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; } // src/vec3.h:79-82
inline double length_squared(V a) { return a.x * a.x + a.y * a.y + a.z * a.z; }  // src/vec3.h:46
inline double length(V a) { return std::sqrt(a.x * a.x + a.y * a.y + a.z * a.z); } // src/vec3.h:45
inline V unit_vector(V a) { return a / length(a); }                             // src/vec3.h:77
inline V vfloor(V a) { return {std::floor(a.x), std::floor(a.y), std::floor(a.z)}; } // src/vec3.h:84
inline double &comp(V &a, int i) { return i == 0 ? a.x : (i == 1 ? a.y : a.z); }

struct Ray { // src/ray.h
  V o, d;
  double tm;
};
inline V at(const Ray &r, double t) { return r.o + t * r.d; } // src/ray.h:11

struct Interval { // src/interval.h
  double min, max;
};
inline bool contains(Interval i, double x) { return i.min <= x && x <= i.max; } // src/interval.h:14

struct Rec { // src/hittable.h:7-30
  V p, normal;
  double t, u, v;
  bool front_face;
  int mat;
  int prim;
};
inline void set_face_normal(Rec &rec, const Ray &r, V outward) { // src/hittable.h:26-29
  rec.front_face = dot(r.d, outward) < 0.0;
  rec.normal = rec.front_face ? outward : -outward;
}

// ------------------------------------------------------------------ Philox4x32-10 draw addressing
// (shared with cpu-ray-tracing-implementation_b200/csrc/philox.cuh; written independently there)
enum Purpose { P_CAMERA = 1, P_LENS = 2, P_MATERIAL = 3, P_SAMPLE = 4, P_VOLUME = 5 };
struct Draws {
  double u[4];
};
inline uint32_t mulhi(uint32_t a, uint32_t b) { return (uint32_t)(((uint64_t)a * b) >> 32); }
Draws philox(uint64_t seed, uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3) {
  uint32_t k0 = (uint32_t)seed, k1 = (uint32_t)(seed >> 32);
  for (int r = 0; r < 10; r++) {
    uint32_t h0 = mulhi(0xD2511F53u, c0), l0 = 0xD2511F53u * c0;
    uint32_t h1 = mulhi(0xCD9E8D57u, c2), l1 = 0xCD9E8D57u * c2;
    uint32_t n0 = h1 ^ c1 ^ k0, n1 = l1, n2 = h0 ^ c3 ^ k1, n3 = l0;
    c0 = n0, c1 = n1, c2 = n2, c3 = n3;
    k0 += 0x9E3779B9u;
    k1 += 0xBB67AE85u;
  }
  const double s = 1.0 / 4294967296.0; // [0,1) like src/utility.h:20
  return {{c0 * s, c1 * s, c2 * s, c3 * s}};
}
struct Stream { // identifies the path whose draws are being consumed
  uint64_t seed;
  uint32_t pixel, sample;
  Draws get(int bounce, int purpose, uint32_t aux) const { return philox(seed, pixel, sample, ((uint32_t)bounce << 8) | (uint32_t)purpose, aux); }
};

// ------------------------------------------------------------------ scene access
struct Scene {
  const b200rt_scene_desc *d;
};
struct Ctx {
  Scene sc;
  Stream st;
  int bounce;
  uint64_t rays;
};

// ------------------------------------------------------------------ aabb::hit (src/aabb.h:28-33,45-69)
inline double stdmin(double a, double b) { return (b < a) ? b : a; } // std::min
inline double stdmax(double a, double b) { return (a < b) ? b : a; } // std::max
bool aabb_hit(const double *f, const Ray &r, Interval ray_t) {
  double t1 = (f[0] - r.o.x) / r.d.x, t2 = (f[1] - r.o.x) / r.d.x;
  Interval ix{stdmin(t1, t2), stdmax(t1, t2)};
  t1 = (f[2] - r.o.y) / r.d.y, t2 = (f[3] - r.o.y) / r.d.y;
  Interval iy{stdmin(t1, t2), stdmax(t1, t2)};
  t1 = (f[4] - r.o.z) / r.d.z, t2 = (f[5] - r.o.z) / r.d.z;
  Interval iz{stdmin(t1, t2), stdmax(t1, t2)};
  double t_min = stdmax(stdmax(ix.min, iy.min), iz.min);
  double t_max = stdmin(stdmin(ix.max, iy.max), iz.max);
  if (t_min > ray_t.min) ray_t.min = t_min;
  if (t_max < ray_t.max) ray_t.max = t_max;
  return ray_t.min < ray_t.max;
}

// ------------------------------------------------------------------ sphere (src/sphere.h:40-95)
void get_sphere_uv(V p, double &u, double &v) { // src/sphere.h:90-95
  double theta = std::acos(-p.y);
  double phi = std::atan2(-p.z, p.x) + pi;
  u = phi / (2 * pi);
  v = theta / pi;
}
bool sphere_hit_raw(V center_, double radius, bool moving, V c0, V c1, const Ray &r, Interval iv, Rec &rec) {
  V center = moving ? c0 + r.tm * (c1 - c0) : center_; // src/sphere.h:41-46,83
  double a = dot(r.d, r.d);
  double b = 2.0 * dot(r.d, (r.o - center));
  double c = dot(r.o - center, r.o - center) - radius * radius;
  double discriminant = b * b - 4 * a * c;
  if (discriminant < 0) return false;
  double sqrtd = std::sqrt(discriminant);
  double root = (-b - sqrtd) / (2.0 * a);
  if (!contains(iv, root)) {
    root = (-b + sqrtd) / (2.0 * a);
    if (!contains(iv, root)) return false;
  }
  rec.t = root;
  rec.p = at(r, root);
  V outward = (rec.p - center_) / radius; // center_, not center: src/sphere.h:69
  get_sphere_uv(outward, rec.u, rec.v);
  set_face_normal(rec, r, outward);
  return true;
}
bool sphere_hit(const Scene &s, int row, const Ray &r, Interval iv, Rec &rec) {
  const double *q = s.d->spheres + (size_t)row * B200RT_SPHERE_STRIDE;
  if (!sphere_hit_raw(v3(q[0], q[1], q[2]), q[3], q[4] != 0.0, v3(q[5], q[6], q[7]), v3(q[8], q[9], q[10]), r, iv, rec)) return false;
  rec.mat = s.d->sphere_mat[row];
  return true;
}

// ------------------------------------------------------------------ quad (src/quad.h:30-64)
bool quad_hit(const Scene &s, int row, const Ray &r, Interval iv, Rec &rec) {
  const double *q = s.d->quads + (size_t)row * B200RT_QUAD_STRIDE;
  V corner = v3(q[0], q[1], q[2]), u_ = v3(q[3], q[4], q[5]), v_ = v3(q[6], q[7], q[8]), unorm = v3(q[9], q[10], q[11]);
  V w = v3(q[12], q[13], q[14]); // n / dot(n, n), n = cross(u_, v_): src/quad.h:39-40, hoisted by the adaptor with the same operations
  double D = q[15];              // dot(unorm_, corner_): src/quad.h:32
  double t = (D - dot(unorm, r.o)) / dot(unorm, r.d);
  if (!contains(iv, t)) return false;
  V hit_point = at(r, t);
  V p = hit_point - corner;
  double a = dot(w, cross(p, v_));
  double b = dot(w, cross(u_, p));
  Interval unit{0, 1};
  if (!contains(unit, a) || !contains(unit, b)) return false; // is_interior, src/quad.h:58-64
  rec.u = a;
  rec.v = b;
  rec.t = t;
  rec.p = hit_point;
  rec.mat = s.d->quad_mat[row];
  set_face_normal(rec, r, unorm);
  return true;
}

// ------------------------------------------------------------------ triangle (src/triangle.h:8-40)
bool triangle_hit(const Scene &s, int row, const Ray &r, Interval iv, Rec &rec) {
  const double *q = s.d->tris + (size_t)row * B200RT_TRI_STRIDE;
  V p0 = v3(q[0], q[1], q[2]), p1 = v3(q[3], q[4], q[5]), p2 = v3(q[6], q[7], q[8]), normal = v3(q[9], q[10], q[11]);
  V e1 = p1 - p0, e2 = p2 - p0, sv = r.o - p0;
  V s1 = cross(r.d, e2), s2 = cross(sv, e1);
  V bary = v3(dot(s2, e2), dot(s1, sv), dot(s2, r.d)) / dot(s1, e1);
  double t = bary.x, b0 = bary.y, b1 = bary.z;
  if (t < iv.min || t > iv.max) return false; // NaN passes: SURVEY.md A.3
  if (b0 < 0 || b1 < 0 || b0 + b1 > 1) return false;
  rec.t = t;
  rec.p = at(r, t);
  set_face_normal(rec, r, normal);
  rec.mat = s.d->tri_mat[row];
  rec.u = 0; // not written by the reference (src/triangle.h:35-39); the ABI defines 0
  rec.v = 0;
  if (s.d->tri_uv) { // the backend's opt-in (b200rt_scene_desc::tri_uv): vertex texture coordinates weighted with b0, b1
    const double *w = s.d->tri_uv + (size_t)row * 6;
    rec.u = (1 - b0 - b1) * w[0] + b0 * w[2] + b1 * w[4];
    rec.v = (1 - b0 - b1) * w[1] + b0 * w[3] + b1 * w[5];
  }
  return true;
}

// ------------------------------------------------------------------ graph traversal
bool node_hit(Ctx &cx, int idx, const Ray &r, Interval iv, Rec &rec, int visit);

bool volume_hit(Ctx &cx, const b200rt_node &n, const Ray &r, Interval ray_t, Rec &rec, int visit) { // src/volumne.h:18-46
  const b200rt_volume &vol = cx.sc.d->volumes[n.a];
  Rec rec1, rec2;
  if (!node_hit(cx, vol.boundary_node, r, Interval{-infinity, +infinity}, rec1, 0)) return false;
  if (!node_hit(cx, vol.boundary_node, r, Interval{rec1.t + 0.0001, infinity}, rec2, 0)) return false;
  if (rec1.t < ray_t.min) rec1.t = ray_t.min;
  if (rec2.t > ray_t.max) rec2.t = ray_t.max;
  if (rec1.t >= rec2.t) return false;
  if (rec1.t < 0) rec1.t = 0;
  double ray_length = length(r.d);
  double distance_inside_boundary = (rec2.t - rec1.t) * ray_length;
  double rnd = cx.st.get(cx.bounce, P_VOLUME, (uint32_t)n.c * 8u + (uint32_t)visit).u[0];
  double hit_distance = -1.0 / vol.density * std::log(rnd);
  if (hit_distance > distance_inside_boundary) return false;
  rec.t = rec1.t + hit_distance / ray_length;
  rec.p = at(r, rec.t);
  rec.normal = v3(1, 0, 0);
  rec.front_face = true;
  rec.mat = vol.material;
  rec.u = 0; // not written by the reference; the ABI defines 0
  rec.v = 0;
  return true;
}

inline double getc(const V &a, int i) { return i == 0 ? a.x : (i == 1 ? a.y : a.z); }
bool rotate_hit(Ctx &cx, const b200rt_node &n, int i0, int i1, const Ray &r, Interval iv, Rec &rec, int visit) {
  // src/hittable.h:125-149 (x: i0=1,i1=2), :192-216 (y: 0,2), :259-284 (z: 0,1)
  const double s = n.f[0], c = n.f[1];
  V o = r.o, d = r.d;
  comp(o, i0) = c * getc(r.o, i0) - s * getc(r.o, i1);
  comp(o, i1) = s * getc(r.o, i0) + c * getc(r.o, i1);
  comp(d, i0) = c * getc(r.d, i0) - s * getc(r.d, i1);
  comp(d, i1) = s * getc(r.d, i0) + c * getc(r.d, i1);
  Ray rr{o, d, r.tm};
  if (!node_hit(cx, n.a, rr, iv, rec, visit)) return false;
  V p = rec.p, nm = rec.normal;
  comp(p, i0) = c * getc(rec.p, i0) + s * getc(rec.p, i1);
  comp(p, i1) = -s * getc(rec.p, i0) + c * getc(rec.p, i1);
  comp(nm, i0) = c * getc(rec.normal, i0) + s * getc(rec.normal, i1);
  comp(nm, i1) = -s * getc(rec.normal, i0) + c * getc(rec.normal, i1);
  rec.p = p;
  rec.normal = nm;
  return true;
}

bool node_hit(Ctx &cx, int idx, const Ray &r, Interval iv, Rec &rec, int visit) {
  const b200rt_node &n = cx.sc.d->nodes[idx];
  switch (n.kind) {
  case B200RT_NODE_BVH: { // src/bvh_node.h:49-59
    if (!aabb_hit(n.f, r, iv)) return false;
    // a one-object node has left_ == right_ and visits the object twice (:34-36); `visit` numbers those
    // visits (binary path code over such ancestors) so that a volume's free-flight draws stay independent
    const bool twice = n.a == n.b;
    bool hit_left = node_hit(cx, n.a, r, iv, rec, twice ? visit * 2 : visit);
    Interval upd{iv.min, hit_left ? rec.t : iv.max};
    bool hit_right = node_hit(cx, n.b, r, upd, rec, twice ? visit * 2 + 1 : visit);
    return hit_left || hit_right;
  }
  case B200RT_NODE_LIST: { // src/hittable_list.h:20-31
    Rec tmp;
    bool any = false;
    for (int k = 0; k < n.b; k++) {
      if (node_hit(cx, (int)cx.sc.d->list_children[n.a + k], r, iv, tmp, visit)) {
        any = true;
        iv.max = tmp.t;
        rec = tmp;
      }
    }
    return any;
  }
  case B200RT_NODE_TRANSLATE: { // src/hittable.h:75-82
    V off = v3(n.f[0], n.f[1], n.f[2]);
    Ray rr{r.o - off, r.d, r.tm};
    if (!node_hit(cx, n.a, rr, iv, rec, visit)) return false;
    rec.p = rec.p + off;
    return true;
  }
  case B200RT_NODE_ROTATE_X: return rotate_hit(cx, n, 1, 2, r, iv, rec, visit);
  case B200RT_NODE_ROTATE_Y: return rotate_hit(cx, n, 0, 2, r, iv, rec, visit);
  case B200RT_NODE_ROTATE_Z: return rotate_hit(cx, n, 0, 1, r, iv, rec, visit);
  case B200RT_NODE_SPHERE:
    if (!sphere_hit(cx.sc, n.a, r, iv, rec)) return false;
    rec.prim = n.c;
    return true;
  case B200RT_NODE_QUAD:
    if (!quad_hit(cx.sc, n.a, r, iv, rec)) return false;
    rec.prim = n.c;
    return true;
  case B200RT_NODE_TRIANGLE:
    if (!triangle_hit(cx.sc, n.a, r, iv, rec)) return false;
    rec.prim = n.c;
    return true;
  case B200RT_NODE_VOLUME:
    if (!volume_hit(cx, n, r, iv, rec, visit)) return false;
    rec.prim = n.c;
    return true;
  }
  return false;
}

// ------------------------------------------------------------------ noise (src/noise.h)
double perlin_interp(const double *vec, const int32_t *perm, int iu, int iv, int iw, double u, double v, double w) { // src/noise.h:22-43
  iu &= 255;
  iv &= 255;
  iw &= 255;
  double uu = u * u * (3 - 2 * u), vv = v * v * (3 - 2 * v), ww = w * w * (3 - 2 * w);
  double accum = 0.0;
  for (int i = 0; i < 2; i++)
    for (int j = 0; j < 2; j++)
      for (int k = 0; k < 2; k++) {
        V weight_v = v3(u - i, v - j, w - k);
        int h = perm[(iu + i) % 256] ^ perm[(iv + j) % 256] ^ perm[(iw + k) % 256]; // perm_x three times: :36
        V g = v3(vec[3 * h], vec[3 * h + 1], vec[3 * h + 2]);
        accum += (i * uu + (1 - i) * (1 - uu)) * (j * vv + (1 - j) * (1 - vv)) * (k * ww + (1 - k) * (1 - ww)) * dot(g, weight_v);
      }
  return accum;
}
double perlin_noise(const double *vec, const int32_t *perm, V p) { // src/noise.h:57-69
  int iu = int(std::floor(p.x)), iv = int(std::floor(p.y)), iw = int(std::floor(p.z));
  double du = p.x - iu, dv = p.y - iv, dw = p.z - iw;
  return perlin_interp(vec, perm, iu, iv, iw, du, dv, dw);
}
double perlin_turb(const double *vec, const int32_t *perm, int depth, V p) { // src/noise.h:45-55
  double accum = 0;
  V temp_p = p;
  double weight = 1.0;
  for (int i = 0; i < depth; i++) {
    accum += weight * perlin_noise(vec, perm, temp_p);
    weight *= 0.5;
    temp_p = temp_p * 2.0;
  }
  return std::fabs(accum);
}
inline double lerp(double t, double a, double b) { return (1 - t) * a + t * b; } // src/utility.h:84
double value_noise(const float *lat, int res, V p) { // src/noise.h:106-132
  V fp = vfloor(p);
  double R = res;
  auto at_ = [&](double idx) -> float { return lat[(size_t)idx]; };
  float x0y0z0 = at_(fp.x * R * R + fp.y * R + fp.z);
  float x1y0z0 = at_((fp.x + 1) * R * R + fp.y * R + fp.z);
  float x0y1z0 = at_(fp.x * R * R + (fp.y + 1) * R + fp.z);
  float x1y1z0 = at_((fp.x + 1) * R * R + (fp.y + 1) * R + fp.z);
  float x0y0z1 = at_(fp.x * R * R + fp.y * R + fp.z + 1);
  float x1y0z1 = at_((fp.x + 1) * R * R + fp.y * R + fp.z + 1);
  float x0y1z1 = at_(fp.x * R * R + (fp.y + 1) * R + fp.z + 1);
  float x1y1z1 = at_((fp.x + 1) * R * R + (fp.y + 1) * R + fp.z + 1);
  double x = p.x - fp.x, y = p.y - fp.y, z = p.z - fp.z;
  double y0z0 = lerp(x, x0y0z0, x1y0z0), y1z0 = lerp(x, x0y1z0, x1y1z0);
  double y0z1 = lerp(x, x0y0z1, x1y0z1), y1z1 = lerp(x, x0y1z1, x1y1z1);
  double z0 = lerp(y, y0z0, y1z0), z1 = lerp(y, y0z1, y1z1);
  return lerp(z, z0, z1);
}
V cell_offset(V u) { // worley_noise::get_rand_offset, src/noise.h:141-145 (== voronoi's, :173-177)
  V rv = v3(dot(u, v3(127.1, 311.7, 74.7)), dot(u, v3(269.5, 183.3, 246.1)), dot(u, v3(113.5, 271.9, 307.7)));
  V res = v3(std::sin(rv.x), std::sin(rv.y), std::sin(rv.z)) * 43758.5453;
  return res - vfloor(res);
}
double worley_noise(V p, bool voronoi) { // src/noise.h:147-167 and :179-200
  V fp = vfloor(p);
  float min_dist = std::numeric_limits<float>::max();
  float colr = 0.0f;
  for (int i = -1; i <= 1; i++)
    for (int j = -1; j <= 1; j++)
      for (int k = -1; k <= 1; k++) {
        V cell = fp + v3(i, j, k);
        V pos = cell + cell_offset(cell);
        float dist = (float)length(pos - p);
        if (dist < min_dist) {
          min_dist = dist;
          if (voronoi) colr = (float)cell_offset(pos).x;
        }
      }
  return voronoi ? (double)colr : (double)(min_dist * min_dist);
}

// ------------------------------------------------------------------ textures (src/texture.h)
// 8-bit texels as the reference reads them (byte / 256, src/texture.h:72-73); float texels (the opt-in ingest) as they are
V image_texel(const Scene &s, const b200rt_image &im, int i, int j) {
  const uint8_t *px = s.d->image_bytes + im.offset + ((size_t)j * im.width + i) * im.bytes_per_pixel;
  if (im.bytes_per_pixel >= 12) {
    float f[3];
    std::memcpy(f, px, 12);
    return v3((double)f[0], (double)f[1], (double)f[2]);
  }
  double cs = 1 / 256.0;
  return v3(px[0] * cs, px[1] * cs, px[2] * cs);
}

V tex_sample(const Scene &s, int tex, double u, double v, V p) {
  const b200rt_texture &t = s.d->textures[tex];
  switch (t.kind) {
  case B200RT_TEX_SOLID: return v3(t.c0[0], t.c0[1], t.c0[2]); // :15
  case B200RT_TEX_CHECKER: {                                    // :47-56
    V q = p / t.scale;
    int ix = (int)std::floor(q.x), iy = (int)std::floor(q.y), iz = (int)std::floor(q.z);
    int total = ix + iy + iz;
    return (total % 2 == 0) ? v3(t.c1[0], t.c1[1], t.c1[2]) : v3(t.c0[0], t.c0[1], t.c0[2]);
  }
  case B200RT_TEX_PICTURE: { // :67-74 with image::pixel_data, src/image.h:72-97
    const b200rt_image &im = s.d->images[t.aux];
    if (!im.has_data) return v3(255 * (1 / 256.0), 0 * (1 / 256.0), 255 * (1 / 256.0));
    int i = (int)(im.width * u);
    int j = (int)(im.height * (1 - v));
    auto clampi = [](int x, int low, int high) { return x < low ? low : (x < high ? x : high - 1); };
    i = clampi(i, 0, im.width);
    j = clampi(j, 0, im.height);
    return image_texel(s, im, i, j);
  }
  case B200RT_TEX_PICTURE_LINEAR: { // the backend's own opt-in (not in the reference): bilinear between texel centres, clamped edges
    const b200rt_image &im = s.d->images[t.aux];
    if (!im.has_data) return v3(255 / 256.0, 0.0, 255 / 256.0);
    double x = im.width * u - 0.5, y = im.height * (1 - v) - 0.5;
    double fx0 = std::floor(x), fy0 = std::floor(y);
    double ax = x - fx0, ay = y - fy0;
    auto cl = [](double f, int n) { return f < 0 ? 0 : (f > n - 1 ? n - 1 : (int)f); };
    int i0 = cl(fx0, im.width), i1 = cl(fx0 + 1, im.width), j0 = cl(fy0, im.height), j1 = cl(fy0 + 1, im.height);
    V c00 = image_texel(s, im, i0, j0), c10 = image_texel(s, im, i1, j0), c01 = image_texel(s, im, i0, j1), c11 = image_texel(s, im, i1, j1);
    V top = (1 - ax) * c00 + ax * c10, bot = (1 - ax) * c01 + ax * c11;
    return (1 - ay) * top + ay * bot;
  }
  case B200RT_TEX_PERLIN: { // :86-88
    const double *vec = s.d->perlin_vec + (size_t)t.aux * 768;
    const int32_t *perm = s.d->perlin_perm + (size_t)t.aux * 256;
    return v3(.5, .5, .5) * (1 + std::sin((p.x + 70 * perlin_turb(vec, perm, 7, p / t.scale))));
  }
  case B200RT_TEX_VALUE: { // :99
    const b200rt_value_noise &vn = s.d->value_noise[t.aux];
    double k = value_noise(s.d->value_lattice + vn.offset, vn.resolution, p);
    return v3(k, k, k);
  }
  case B200RT_TEX_WORLEY: {
    double k = worley_noise(p, false); // :107
    return v3(k, k, k);
  }
  case B200RT_TEX_VORONOI: {
    double k = worley_noise(p, true); // :115
    return v3(k, k, k);
  }
  }
  return v3(0, 0, 0);
}

// ------------------------------------------------------------------ sampling helpers (src/utility.h, src/onb.h, src/pdf.h)
V random_in_unit_sphere(double u1, double u2) { // src/utility.h:30-42
  double cos_theta = 1 - 2 * u1;
  double sin_theta = std::sqrt(1 - cos_theta * cos_theta);
  double phi = 2 * pi * u2;
  return v3(sin_theta * std::cos(phi), cos_theta, sin_theta * std::sin(phi));
}
V random_cosine_direction(double r1, double r2) { // src/utility.h:61-69
  double phi = 2 * pi * r1;
  return v3(std::cos(phi) * std::sqrt(r2), std::sqrt(1 - r2), std::sin(phi) * std::sqrt(r2));
}
struct Onb { // src/onb.h:18-28
  V x, y, z;
  explicit Onb(V normal) {
    y = unit_vector(normal);
    V a = (std::fabs(y.x) > 0.9) ? v3(0, 0, 1) : v3(1, 0, 0);
    z = unit_vector(cross(y, a));
    x = cross(y, z);
  }
  V transform(V v) const { return v3(0, 0, 0) + v.x * x + v.y * y + v.z * z; } // src/onb.h:6
};
inline V reflect(V v, V n) { return v - 2 * dot(v, n) * n; } // src/utility.h:70
V refract(V v, V n, double eta) {                            // src/utility.h:71-76
  double cos_theta = std::fmin(dot(-v, n), 1.0);
  V perp = eta * (v + cos_theta * n);
  V par = -std::sqrt(std::fabs(1.0 - length_squared(perp))) * n;
  return perp + par;
}
double cosine_pdf_value(const Onb &uvw, V direction) { // src/pdf.h:37-40
  double c = dot(unit_vector(direction), uvw.y);
  return std::fmax(0, c / pi);
}

// light = the hittable handed to camera::render; hittable_pdf forwards to it (src/hittable_list.h:39-50)
double light_pdf_value(const Scene &s, V origin, V direction) {
  switch (s.d->light_kind) {
  case B200RT_LIGHT_QUAD: { // src/quad.h:66-73
    Rec rec;
    Ray r{origin, direction, 0.0};
    if (!quad_hit(s, s.d->light_index, r, Interval{0.001, infinity}, rec)) return 0;
    double distance_squared = rec.t * rec.t * length_squared(direction);
    double cosine = std::fabs(dot(unit_vector(direction), rec.normal));
    const double *q = s.d->quads + (size_t)s.d->light_index * B200RT_QUAD_STRIDE;
    return distance_squared / (cosine * q[16]);
  }
  case B200RT_LIGHT_SPHERE: { // src/sphere.h:76-78
    const double *q = s.d->spheres + (size_t)s.d->light_index * B200RT_SPHERE_STRIDE;
    return q[3] * q[3] * pi / length_squared(origin - v3(q[0], q[1], q[2]));
  }
  default: return 0.0; // src/hittable.h:39
  }
}
V light_random(const Scene &s, V origin, double u1, double u2) {
  switch (s.d->light_kind) {
  case B200RT_LIGHT_QUAD: { // src/quad.h:75-78
    const double *q = s.d->quads + (size_t)s.d->light_index * B200RT_QUAD_STRIDE;
    V p = v3(q[0], q[1], q[2]) + (u1 * v3(q[3], q[4], q[5])) + (u2 * v3(q[6], q[7], q[8]));
    return p - origin;
  }
  case B200RT_LIGHT_SPHERE: { // src/sphere.h:81
    const double *q = s.d->spheres + (size_t)s.d->light_index * B200RT_SPHERE_STRIDE;
    return random_in_unit_sphere(u1, u2) * q[3];
  }
  default: return v3(1, 0, 0); // src/hittable.h:41
  }
}

// ------------------------------------------------------------------ camera::miss (src/camera.h:180-190)
V miss(const Scene &s, const Ray &r) {
  if (s.d->background_tex < 0) return v3(0, 0, 0);
  Rec rec;
  // sphere unit_sphere(r.origin(), 1.0f, nullptr): center_ = origin, radius_ = fmax(0, 1.0)
  if (sphere_hit_raw(r.o, 1.0, false, v3(0, 0, 0), v3(0, 0, 0), r, Interval{0.001, infinity}, rec)) return tex_sample(s, s.d->background_tex, rec.u, rec.v, rec.p);
  return v3(0, 0, 0);
}

// ------------------------------------------------------------------ one level of ray_color below the hit test
// (src/camera.h:202-240): material::emitted, material::scatter, the pdf the reference builds, generate / value, p_scattered.
// dm = the draws scatter() consumes, ds = the draws the pdf consumes (addressing documented at b200rt_scatter, include/b200rt.h).
struct Shade {
  int mode;      // 0 = scatter() returned false, 1 = kDetermined, 2 = kRandom
  V emitted;     // color_from_emission (:205)
  V attenuation; // scatter_record.attenuation
  V dir;         // direction of the scattered ray
  double pdf_value, p_scattered;
};
Shade shade_step(const Scene &s, const Rec &rec, const Ray &r, const Draws &dm, const Draws &ds) {
  Shade o{0, v3(0, 0, 0), v3(0, 0, 0), v3(0, 0, 0), 0.0, 0.0};
  const b200rt_material &m = s.d->materials[rec.mat];
  if (m.kind == B200RT_MAT_DIFFUSE_LIGHT) { // src/material.h:211-215; scatter() is the base's `false` (:43)
    if (rec.front_face) o.emitted = tex_sample(s, m.tex, rec.u, rec.v, rec.p);
    return o;
  }
  int pdf_kind = 0; // 1 cosine, 2 sphere
  switch (m.kind) {
  case B200RT_MAT_LAMBERTIAN: // src/material.h:62-67
    o.attenuation = tex_sample(s, m.tex, rec.u, rec.v, rec.p);
    pdf_kind = 1;
    break;
  case B200RT_MAT_METAL: { // src/material.h:85-92
    V dir = unit_vector(reflect(r.d, rec.normal));
    dir = dir + (double)m.p0 * unit_vector(random_in_unit_sphere(dm.u[1], dm.u[2]));
    o.mode = 1;
    o.dir = dir;
    o.attenuation = tex_sample(s, m.tex, rec.u, rec.v, rec.p);
    break;
  }
  case B200RT_MAT_DIELECTRIC: { // src/material.h:113-131
    o.mode = 1;
    o.attenuation = tex_sample(s, m.tex, rec.u, rec.v, rec.p);
    double ri = rec.front_face ? (1.0 / m.p0) : m.p0;
    V ud = unit_vector(r.d);
    double cos_theta = std::fmin(dot(-ud, rec.normal), 1.0);
    double sin_theta = std::sqrt(1.0 - cos_theta * cos_theta);
    bool cant_refract = ri * sin_theta > 1.0;
    auto reflectance = [](double cosine, double ri) { // :135-139
      double r0 = (1 - ri) / (1 + ri);
      r0 = r0 * r0;
      return r0 + (1 - r0) * std::pow((1 - cosine), 5);
    };
    if (cant_refract || reflectance(cos_theta, ri) > dm.u[0])
      o.dir = reflect(ud, rec.normal);
    else
      o.dir = refract(ud, rec.normal, ri);
    break;
  }
  case B200RT_MAT_GLOSS: { // src/material.h:158-174
    V specular_dir = reflect(r.d, rec.normal);
    bool is_specular = dm.u[0] <= (double)m.p1;
    if (is_specular) {
      V diffuse_dir = Onb(rec.normal).transform(random_cosine_direction(dm.u[1], dm.u[2]));
      double t = (double)m.p0;
      o.dir = unit_vector((1 - t) * diffuse_dir + t * specular_dir); // linear_interpolation, src/utility.h:85
      o.mode = 1;
      o.attenuation = v3(1.0, 1.0, 1.0);
    } else {
      o.attenuation = tex_sample(s, m.tex, rec.u, rec.v, rec.p);
      pdf_kind = 1;
    }
    break;
  }
  case B200RT_MAT_ISOTROPIC: // src/material.h:193-198
    o.attenuation = tex_sample(s, m.tex, rec.u, rec.v, rec.p);
    pdf_kind = 2;
    break;
  }
  if (o.mode == 1) return o; // kDetermined, src/camera.h:210-214

  // kRandom: src/camera.h:217-240
  o.mode = 2;
  Onb uvw(pdf_kind == 1 ? rec.normal : v3(0, 1, 0));
  auto mat_generate = [&]() -> V {
    if (pdf_kind == 1) return uvw.transform(random_cosine_direction(ds.u[1], ds.u[2])); // src/pdf.h:41
    return unit_vector(random_in_unit_sphere(ds.u[1], ds.u[2]));                        // src/pdf.h:19
  };
  auto mat_value = [&](V dir) -> double { return pdf_kind == 1 ? cosine_pdf_value(uvw, dir) : 1 / (4 * pi); };
  if (s.d->light_kind == B200RT_LIGHT_NONE) {
    o.dir = mat_generate();
    o.pdf_value = mat_value(o.dir);
  } else { // dual_pdf, src/pdf.h:48-61
    o.dir = (ds.u[0] < 0.5) ? light_random(s, rec.p, ds.u[1], ds.u[2]) : mat_generate();
    o.pdf_value = 0.5 * light_pdf_value(s, rec.p, o.dir) + 0.5 * mat_value(o.dir);
  }
  if (pdf_kind == 1) { // lambertian / gloss p_scattered, src/material.h:69-72,176-179
    double cosine = dot(rec.normal, unit_vector(o.dir));
    o.p_scattered = cosine < 0 ? 0 : cosine / pi;
  } else {
    o.p_scattered = 1 / (4 * pi); // src/material.h:200
  }
  return o;
}

// ------------------------------------------------------------------ camera::ray_color (src/camera.h:193-241)
V ray_color(Ctx &cx, const Ray &r, int iteration) {
  if (iteration <= 0) return v3(0, 0, 0);
  const Scene &s = cx.sc;
  Rec rec;
  const V &d = r.d;
  if (!(std::isnan(d.x) || std::isnan(d.y) || std::isnan(d.z))) cx.rays++; // counting rule of SURVEY.md §8d
  if (!node_hit(cx, (int)s.d->root, r, Interval{0.001, infinity}, rec, 0)) return miss(s, r);

  const int b = cx.bounce;
  const Shade sh = shade_step(s, rec, r, cx.st.get(b, P_MATERIAL, 0), cx.st.get(b, P_SAMPLE, 0));
  if (sh.mode == 0) return sh.emitted; // :207-208
  cx.bounce = b + 1;
  const Ray scattered{rec.p, sh.dir, r.tm};
  V in = ray_color(cx, scattered, iteration - 1);
  if (sh.mode == 1) return sh.attenuation * in + sh.emitted; // :210-214
  V from_scatter = (sh.attenuation * sh.p_scattered * in) / sh.pdf_value; // :224, :237
  return from_scatter + sh.emitted;
}

// ------------------------------------------------------------------ camera::generate_ray (src/camera.h:244-284)
// jx, jy: the two sample_square draws (:293); tm: the ray-time draw; disk(it, a, b): the it-th pair of
// random_in_unit_disk draws (src/utility.h:46-52).
template <class DiskDraw> Ray generate_ray_from(const b200rt_camera &c, int y, int x, double jx, double jy, double tm, DiskDraw disk) {
  V pos = v3(c.pos[0], c.pos[1], c.pos[2]), dir = v3(c.dir[0], c.dir[1], c.dir[2]);
  V right = v3(c.right[0], c.right[1], c.right[2]), up = v3(c.up[0], c.up[1], c.up[2]);
  V delta_u = c.viewport_width * right / c.image_width;    // src/camera.h:137
  V delta_v = -c.viewport_height * up / c.image_height;   // src/camera.h:138
  V offset = v3(jx - 0.5, jy - 0.5, 0); // sample_square, :293
  if (c.mode == B200RT_CAM_PERSPECTIVE) {
    V d00 = c.focal_length * dir - c.viewport_width / 2.0 * right + c.viewport_height / 2.0 * up + 0.5 * (delta_u + delta_v);
    V rd = d00 + x * delta_u + y * delta_v;
    V rrd = rd + offset.x * delta_u + offset.y * delta_v;
    return Ray{pos, rrd, tm};
  } else if (c.mode == B200RT_CAM_ORTHOGRAPHIC) {
    V p00 = pos - c.viewport_width / 2.0 * right + c.viewport_height / 2.0 * up + 0.5 * (delta_u + delta_v);
    V p = p00 + x * delta_u + y * delta_v;
    V rp = p + offset.x * delta_u + offset.y * delta_v;
    return Ray{rp, dir, tm};
  } else if (c.mode == B200RT_CAM_FISHEYE) {
    V d00 = c.focal_length * dir - c.viewport_width / 2.0 * right + c.viewport_height / 2.0 * up + 0.5 * (delta_u + delta_v);
    V rd = d00 + x * delta_u + y * delta_v;
    V rrd = rd + offset.x * delta_u + offset.y * delta_v;
    double rr = length(rrd - dir);
    double theta = std::asin(rr / c.focal_length);
    V v1 = unit_vector(dir), v2 = unit_vector(rrd);
    double b_prime = std::sqrt(std::sin(theta) * std::sin(theta) / (1 - dot(v1, v2) * dot(v1, v2)));
    double a_prime = std::cos(theta) - b_prime * dot(v1, v2);
    V v3_ = a_prime * v1 + b_prime * v2;
    return Ray{pos, v3_, tm};
  } else { // lens, :276-283
    V f00 = pos - c.viewport_width / 2.0 * right + c.viewport_height / 2.0 * up + 0.5 * (delta_u + delta_v);
    V rd = f00 + x * delta_u + y * delta_v + offset.x * delta_u + offset.y * delta_v + c.focus_dist * dir;
    V p;
    for (uint32_t it = 0;; it++) { // random_in_unit_disk, src/utility.h:46-52
      double a, b;
      disk(it, a, b);
      p = v3(-1 + (1 - -1) * a, -1 + (1 - -1) * b, 0); // random_double(-1, 1), src/utility.h:22
      if (length_squared(p) < 1) break;
    }
    V du = v3(c.defocus_disk_u[0], c.defocus_disk_u[1], c.defocus_disk_u[2]);
    V dv = v3(c.defocus_disk_v[0], c.defocus_disk_v[1], c.defocus_disk_v[2]);
    V ro = pos + ((p.x * du) + (p.y * dv)); // sample_defocus_disk, :287-290
    rd = rd - ro;
    return Ray{ro, rd, 0.0};
  }
}
Ray generate_ray(const b200rt_camera &c, const Stream &st, int y, int x) {
  Draws dr = st.get(0, P_CAMERA, 0);
  return generate_ray_from(c, y, x, dr.u[0], dr.u[1], dr.u[2], [&](uint32_t it, double &a, double &b) {
    Draws dl = st.get(0, P_LENS, it);
    a = dl.u[0];
    b = dl.u[1];
  });
}

} // namespace

extern "C" {

struct oracle_stats {
  uint64_t samples, rays, nan_samples;
  double seconds;
};

int oracle_trace(const b200rt_scene_desc *d, const double *rays7, size_t n, double tmin, double tmax, uint64_t seed, int32_t *prim, double *t,
                 double *uv, uint8_t *front, double *p, double *normal) {
  // rays are independent (the volume draw of ray i is keyed by i): large batches are spread over the host threads
  const unsigned hw = std::max(1u, std::thread::hardware_concurrency());
  const size_t n_thr = n >= 65536 ? hw : 1;
  auto work = [&](size_t lo, size_t hi) {
  for (size_t i = lo; i < hi; i++) {
    const double *q = rays7 + 7 * i;
    Ray r{v3(q[0], q[1], q[2]), v3(q[3], q[4], q[5]), q[6]};
    Ctx cx{Scene{d}, Stream{seed, (uint32_t)i, (uint32_t)(i >> 32)}, 0, 0};
    Rec rec;
    std::memset(&rec, 0, sizeof rec);
    bool ok = node_hit(cx, (int)d->root, r, Interval{tmin, tmax}, rec, 0);
    if (prim) prim[i] = ok ? rec.prim : -1;
    if (t) t[i] = ok ? rec.t : 0.0;
    if (uv) uv[2 * i] = ok ? rec.u : 0.0, uv[2 * i + 1] = ok ? rec.v : 0.0;
    if (front) front[i] = ok && rec.front_face ? 1 : 0;
    const double pp[3] = {rec.p.x, rec.p.y, rec.p.z}, nn[3] = {rec.normal.x, rec.normal.y, rec.normal.z};
    for (int c = 0; c < 3; c++) {
      if (p) p[3 * i + c] = ok ? pp[c] : 0.0;
      if (normal) normal[3 * i + c] = ok ? nn[c] : 0.0;
    }
  }
  };
  if (n_thr == 1) {
    work(0, n);
  } else {
    std::vector<std::thread> pool;
    for (size_t k = 0; k < n_thr; k++) pool.emplace_back(work, n * k / n_thr, n * (k + 1) / n_thr);
    for (auto &th : pool) th.join();
  }
  return 0;
}

int oracle_generate_rays(const b200rt_camera *c, uint64_t seed, int spp_begin, int spp_end, double *rays7) {
  const size_t npix = (size_t)c->image_width * c->image_height;
  for (int k = spp_begin; k < spp_end; k++)
    for (int i = 0; i < c->image_height; i++)
      for (int j = 0; j < c->image_width; j++) {
        uint32_t pix = (uint32_t)(i * c->image_width + j);
        Ray r = generate_ray(*c, Stream{seed, pix, (uint32_t)k}, i, j);
        double *o = rays7 + 7 * ((size_t)(k - spp_begin) * npix + pix);
        o[0] = r.o.x, o[1] = r.o.y, o[2] = r.o.z, o[3] = r.d.x, o[4] = r.d.y, o[5] = r.d.z, o[6] = r.tm;
      }
  return 0;
}

// generate_ray on caller-supplied draws (the values the reference consumed, in its consumption order).
// GCC evaluates the two arguments of `vec3(random_double() - 0.5, random_double() - 0.5, 0)` right to left, so
// with swap_pairs != 0 the FIRST draw of each pair feeds y and the second x.  Draw order: jitter pair,
// then the time draw (perspective / orthographic / fisheye) or the disk pairs (lens).
int oracle_generate_rays_from_draws(const b200rt_camera *c, size_t n, const double *draws, int max_draws, int swap_pairs, double *rays7) {
  const size_t npix = (size_t)c->image_width * c->image_height;
  for (size_t idx = 0; idx < n; idx++) {
    const double *u = draws + idx * max_draws;
    size_t pix = idx % npix;
    int i = (int)(pix / c->image_width), j = (int)(pix % c->image_width);
    double jx = swap_pairs ? u[1] : u[0], jy = swap_pairs ? u[0] : u[1];
    Ray r = generate_ray_from(*c, i, j, jx, jy, u[2], [&](uint32_t it, double &a, double &b) {
      const double *q = u + 2 + 2 * it;
      a = swap_pairs ? q[1] : q[0];
      b = swap_pairs ? q[0] : q[1];
    });
    double *o = rays7 + 7 * idx;
    o[0] = r.o.x, o[1] = r.o.y, o[2] = r.o.z, o[3] = r.d.x, o[4] = r.d.y, o[5] = r.d.z, o[6] = r.tm;
  }
  return 0;
}

int oracle_texture_sample(const b200rt_scene_desc *d, int tex, const double *uvp5, size_t n, double *rgb) {
  Scene s{d};
  for (size_t i = 0; i < n; i++) {
    const double *q = uvp5 + 5 * i;
    V c = tex_sample(s, tex, q[0], q[1], v3(q[2], q[3], q[4]));
    rgb[3 * i] = c.x, rgb[3 * i + 1] = c.y, rgb[3 * i + 2] = c.z;
  }
  return 0;
}

// ray_color on caller-supplied rays; ray i consumes the draws of stream (pixel = i, sample = 0).
// b200rt_scatter's counterpart: same buffers (include/b200rt.h).
int oracle_scatter(const b200rt_scene_desc *d, const double *rays7, const double *rec12, const double *draws6, size_t n, double *out16) {
  Scene s{d};
  for (size_t i = 0; i < n; i++) {
    const double *q = rays7 + 7 * i, *h = rec12 + 12 * i, *u = draws6 + 6 * i;
    Ray r{v3(q[0], q[1], q[2]), v3(q[3], q[4], q[5]), q[6]};
    Rec rec{};
    rec.p = v3(h[0], h[1], h[2]), rec.normal = v3(h[3], h[4], h[5]);
    rec.t = h[6], rec.u = h[7], rec.v = h[8];
    rec.front_face = h[9] != 0.0;
    rec.mat = (int)h[10];
    if (rec.mat < 0 || (uint32_t)rec.mat >= d->n_materials) return -1;
    const Shade sh = shade_step(s, rec, r, Draws{{u[0], u[1], u[2], 0.0}}, Draws{{u[3], u[4], u[5], 0.0}});
    double *o = out16 + 16 * i;
    o[0] = sh.mode;
    o[1] = sh.emitted.x, o[2] = sh.emitted.y, o[3] = sh.emitted.z;
    o[4] = sh.attenuation.x, o[5] = sh.attenuation.y, o[6] = sh.attenuation.z;
    o[7] = sh.dir.x, o[8] = sh.dir.y, o[9] = sh.dir.z;
    o[10] = sh.pdf_value, o[11] = sh.p_scattered;
    V w = sh.mode == 0 ? v3(0, 0, 0) : (sh.mode == 1 ? sh.attenuation : (sh.attenuation * sh.p_scattered) / sh.pdf_value);
    o[12] = w.x, o[13] = w.y, o[14] = w.z, o[15] = 0.0;
  }
  return 0;
}

int oracle_ray_color(const b200rt_scene_desc *d, const double *rays7, size_t n, int depth, uint64_t seed, double *rgb) {
  for (size_t i = 0; i < n; i++) {
    const double *q = rays7 + 7 * i;
    Ray r{v3(q[0], q[1], q[2]), v3(q[3], q[4], q[5]), q[6]};
    Ctx cx{Scene{d}, Stream{seed, (uint32_t)i, 0u}, 0, 0};
    V c = ray_color(cx, r, depth);
    rgb[3 * i] = c.x, rgb[3 * i + 1] = c.y, rgb[3 * i + 2] = c.z;
  }
  return 0;
}

// The sample loop of camera::render (src/camera.h:163-170) for samples [spp_begin, spp_end) of every pixel.
// rgb_sum gets the per-pixel SUM in sample order (what b200rt_render returns); rgb_sumsq (nullable) the sum of
// squares over finite samples.  per_sample (nullable) receives every sample's radiance, index
// ((k - spp_begin) * W*H + pixel) * 3.
int oracle_render(const b200rt_scene_desc *d, const b200rt_camera *c, uint64_t seed, int spp_begin, int spp_end, int threads, double *rgb_sum,
                  double *rgb_sumsq, double *per_sample, oracle_stats *st) {
  const int W = c->image_width, H = c->image_height;
  const size_t npix = (size_t)W * H;
  if (threads < 1) threads = 1;
  std::atomic<int> next_row{0};
  std::atomic<uint64_t> rays{0}, nans{0};
  auto worker = [&]() {
    uint64_t my_rays = 0, my_nans = 0;
    for (;;) {
      int i = next_row.fetch_add(1);
      if (i >= H) break;
      for (int j = 0; j < W; j++) {
        uint32_t pix = (uint32_t)(i * W + j);
        double a[3] = {0, 0, 0}, b[3] = {0, 0, 0};
        for (int k = spp_begin; k < spp_end; k++) {
          Ctx cx{Scene{d}, Stream{seed, pix, (uint32_t)k}, 0, 0};
          Ray r = generate_ray(*c, cx.st, i, j);
          V col;
          if (std::isnan(r.d.x) || std::isnan(r.d.y) || std::isnan(r.d.z)) {
            // fisheye outside the image circle: the reference's NaN ray contributes exactly 0
            col = v3(0, 0, 0);
            my_nans++;
          } else {
            col = ray_color(cx, r, c->max_depth);
          }
          my_rays += cx.rays;
          const double cc[3] = {col.x, col.y, col.z};
          for (int q = 0; q < 3; q++) {
            a[q] += cc[q];
            if (std::isfinite(cc[q])) b[q] += cc[q] * cc[q];
            if (per_sample) per_sample[((size_t)(k - spp_begin) * npix + pix) * 3 + q] = cc[q];
          }
        }
        for (int q = 0; q < 3; q++) {
          if (rgb_sum) rgb_sum[3 * (size_t)pix + q] = a[q];
          if (rgb_sumsq) rgb_sumsq[3 * (size_t)pix + q] = b[q];
        }
      }
    }
    rays += my_rays;
    nans += my_nans;
  };
  auto t0 = std::chrono::steady_clock::now();
  if (threads == 1)
    worker();
  else {
    std::vector<std::thread> pool;
    for (int t = 0; t < threads; t++) pool.emplace_back(worker);
    for (auto &t : pool) t.join();
  }
  if (st) {
    st->samples = npix * (uint64_t)(spp_end - spp_begin);
    st->rays = rays.load();
    st->nan_samples = nans.load();
    st->seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
  }
  return 0;
}

} // extern "C"
