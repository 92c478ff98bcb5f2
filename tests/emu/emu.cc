// emu.cc — TEST INFRASTRUCTURE.  Runs the device functions of libb200rt.so's source (flatten.h, bvh_build.h,
// traverse.cuh, mirror.cuh, intersect.cuh, shading.cuh, camera.cuh) on the CPU, one path at a time, so that the
// CPU-only test tier can check the flattener, the BVH builder, the tie-break rule and the shading logic against
// the oracle before any GPU time is spent.  It is NOT a fallback: nothing outside tests/ loads it, and the
// wavefront kernels themselves (queues, compaction, atomics) are only exercised on a GPU.
#include "bvh_build.h"
#include "camera.cuh"
#include "mirror.cuh"
#include "shading.cuh"
#include "traverse.cuh"

using namespace b200;

namespace {
struct EmuScene {
  DevScene S;
  FlatScene fs;
  BuiltBvh bvh;
  std::vector<Entry> ordered;
};
} // namespace

extern "C" {

void *emu_scene_create(const b200rt_scene_desc *d, int leaf_size) {
  auto *e = new EmuScene;
  try {
    Flattener fl(*d);
    fl.run();
    e->fs = std::move(fl.out);
    if (leaf_size >= 100) { // the host builder on ONE thread (leaf_size - 100 per leaf): the tree the threaded build has to reproduce
      e->bvh = BvhBuilder(e->fs.boxes, leaf_size - 100, true, false, false, 1).run();
    } else if (leaf_size >= 0) {
      e->bvh = BvhBuilder(e->fs.boxes, leaf_size).run();
    } else if (leaf_size <= -10) { // the library's small-scene alternative (bvh_build.h: build_agglomerative), -leaf_size - 10 per leaf
      e->bvh = build_agglomerative(e->fs.boxes, -leaf_size - 10);
    } else {
      // the device PLOC build (csrc/bvh_ploc.cuh) restated serially: clusters in flattener order (the device sorts by Morton code
      // first; any order gives a valid tree), rounds of mutual nearest neighbours within +-radius, children before parents, the
      // root last; then the host pass the library runs on the device's topology (bvh_build.h:finish_tree), which is what this
      // mode is for.  -leaf_size = primitives per leaf.
      struct Cl {
        float v[6];
        int32_t ref;
      };
      const std::vector<Box> &bx = e->fs.boxes;
      std::vector<Cl> cl(bx.size());
      double bound[3] = {0, 0, 0};
      for (size_t i = 0; i < bx.size(); i++) {
        for (int k = 0; k < 3; k++) {
          float lo = (float)bx[i].lo[k], hi = (float)bx[i].hi[k];
          if ((double)lo > bx[i].lo[k]) lo = std::nextafterf(lo, -INFINITY);
          if ((double)hi < bx[i].hi[k]) hi = std::nextafterf(hi, INFINITY);
          cl[i].v[2 * k] = lo, cl[i].v[2 * k + 1] = hi;
          for (float f : {lo, hi})
            if (std::isfinite(f) && std::fabs((double)f) > bound[k]) bound[k] = std::fabs((double)f);
        }
        cl[i].ref = ~(int32_t)((uint32_t)i << 3);
      }
      auto union_area = [](const Cl &a, const Cl &b) {
        double d[3];
        for (int k = 0; k < 3; k++) d[k] = (double)std::max(a.v[2 * k + 1], b.v[2 * k + 1]) - (double)std::min(a.v[2 * k], b.v[2 * k]);
        if (!(d[0] >= 0 && d[1] >= 0 && d[2] >= 0)) return 0.0;
        const double ar = 2 * (d[0] * d[1] + d[1] * d[2] + d[2] * d[0]);
        return std::isfinite(ar) ? ar : 1e300;
      };
      std::vector<BvhNode> topo;
      const int radius = 4;
      while (cl.size() > 1) {
        const int m = (int)cl.size();
        std::vector<int> nn(m, -1);
        for (int i = 0; i < m; i++) {
          double best = 0;
          for (int j = std::max(0, i - radius); j <= std::min(m - 1, i + radius); j++) {
            if (j == i) continue;
            const double ar = union_area(cl[i], cl[j]);
            if (nn[i] < 0 || ar < best) best = ar, nn[i] = j;
          }
        }
        std::vector<Cl> next;
        for (int i = 0; i < m; i++) {
          const int j = nn[i];
          const bool mutual = nn[j] == i;
          if (mutual && i > j) continue;
          Cl c = cl[i];
          if (mutual) {
            BvhNode nd;
            std::memset(&nd, 0, sizeof nd);
            for (int k = 0; k < 6; k++) nd.lbox[k] = cl[i].v[k], nd.rbox[k] = cl[j].v[k];
            nd.left = cl[i].ref, nd.right = cl[j].ref;
            for (int k = 0; k < 3; k++) c.v[2 * k] = std::min(c.v[2 * k], cl[j].v[2 * k]), c.v[2 * k + 1] = std::max(c.v[2 * k + 1], cl[j].v[2 * k + 1]);
            c.ref = (int32_t)topo.size();
            topo.push_back(nd);
          }
          next.push_back(c);
        }
        if (next.size() == cl.size()) throw std::runtime_error("no mutual pair");
        cl.swap(next);
      }
      e->bvh = finish_tree(topo, (uint32_t)bx.size(), -leaf_size, bound);
    }
  } catch (const std::exception &ex) {
    delete e;
    return nullptr;
  }
  e->ordered.resize(e->fs.entries.size());
  for (size_t k = 0; k < e->ordered.size(); k++) e->ordered[k] = e->fs.entries[e->bvh.perm[k]];
  DevScene &S = e->S;
  std::memset(&S, 0, sizeof S);
  S.nodes = d->nodes, S.list_children = d->list_children, S.root = d->root;
  S.spheres = d->spheres, S.sphere_mat = d->sphere_mat, S.quads = d->quads, S.quad_mat = d->quad_mat, S.tris = d->tris, S.tri_mat = d->tri_mat;
  S.volumes = d->volumes, S.materials = d->materials, S.textures = d->textures, S.images = d->images, S.image_bytes = d->image_bytes;
  S.perlin_vec = d->perlin_vec, S.perlin_perm = d->perlin_perm, S.value_noise = d->value_noise, S.value_lattice = d->value_lattice;
  S.light_kind = d->light_kind, S.light_index = d->light_index, S.background_tex = d->background_tex;
  S.bvh = e->bvh.nodes.data(), S.bvh_root = e->bvh.root, S.n_bvh_nodes = (uint32_t)e->bvh.nodes.size();
  S.entries = e->ordered.data(), S.bentries = e->fs.bentries.data(), S.vol_inst = e->fs.vol_inst.data();
  S.chain_ops = e->fs.chain_ops.data(), S.chains = e->fs.chains.data(), S.n_entries = (uint32_t)e->ordered.size();
  S.has_volumes = e->fs.has_volumes;
  for (int k = 0; k < 3; k++) S.bound[k] = e->bvh.bound[k];
  S.self = &e->S;
  return e;
}
void emu_scene_destroy(void *h) { delete static_cast<EmuScene *>(h); }
uint64_t emu_scene_bvh_hash(void *h) { // FNV-1a over the node array and the entry permutation
  auto *e = static_cast<EmuScene *>(h);
  uint64_t x = 1469598103934665603ull;
  auto eat = [&](const void *p, size_t n) {
    const unsigned char *b = static_cast<const unsigned char *>(p);
    for (size_t i = 0; i < n; i++) x = (x ^ b[i]) * 1099511628211ull;
  };
  eat(e->bvh.nodes.data(), e->bvh.nodes.size() * sizeof(BvhNode));
  eat(e->bvh.perm.data(), e->bvh.perm.size() * sizeof(uint32_t));
  eat(&e->bvh.root, sizeof e->bvh.root);
  return x;
}
void emu_scene_info(void *h, uint32_t *out6) {
  auto *e = static_cast<EmuScene *>(h);
  out6[0] = (uint32_t)e->bvh.nodes.size(), out6[1] = e->bvh.n_leaves, out6[2] = (uint32_t)e->ordered.size(), out6[3] = e->bvh.max_depth;
  out6[4] = (uint32_t)e->fs.graph_depth, out6[5] = (uint32_t)e->fs.chains.size();
}

int emu_trace(void *h, int mirror, const double *rays7, size_t n, double lo, double hi, uint64_t seed, int32_t *prim, double *t, double *uv,
              uint8_t *front, double *p, double *normal) {
  const DevScene &S = static_cast<EmuScene *>(h)->S;
  for (size_t i = 0; i < n; i++) {
    const double *q = rays7 + 7 * i;
    const Ray r{v3(q[0], q[1], q[2]), v3(q[3], q[4], q[5]), q[6]};
    const Stream st{seed, (uint32_t)i, (uint32_t)(i >> 32)};
    Rec rec;
    std::memset(&rec, 0, sizeof rec);
    bool ok;
    if (mirror) {
      MirrorCtx cx{&S, st, 0};
      ok = mirror_hit(cx, (int)S.root, r, lo, hi, rec, 0);
    } else {
      const Best b = trace_closest(S, r, lo, hi, st, 0);
      ok = b.entry >= 0;
      if (ok) finish_hit(S, S.entries[b.entry], r, b.t, rec);
    }
    if (prim) prim[i] = ok ? rec.prim : -1;
    if (t) t[i] = ok ? rec.t : 0.0;
    if (uv) uv[2 * i] = ok ? rec.u : 0.0, uv[2 * i + 1] = ok ? rec.v : 0.0;
    if (front) front[i] = ok && rec.front_face ? 1 : 0;
    if (p) p[3 * i] = ok ? rec.p.x : 0.0, p[3 * i + 1] = ok ? rec.p.y : 0.0, p[3 * i + 2] = ok ? rec.p.z : 0.0;
    if (normal) normal[3 * i] = ok ? rec.normal.x : 0.0, normal[3 * i + 1] = ok ? rec.normal.y : 0.0, normal[3 * i + 2] = ok ? rec.normal.z : 0.0;
  }
  return 0;
}

int emu_generate_rays(const b200rt_camera *c, uint64_t seed, int spp_begin, int spp_end, double *rays7) {
  const DevCamera cam = make_camera(*c);
  const size_t npix = (size_t)c->image_width * c->image_height;
  for (size_t i = 0; i < npix * (size_t)(spp_end - spp_begin); i++) {
    const uint32_t pixel = (uint32_t)(i % npix), sample = (uint32_t)spp_begin + (uint32_t)(i / npix);
    const Ray r = generate_ray(cam, Stream{seed, pixel, sample}, (int)(pixel / c->image_width), (int)(pixel % c->image_width));
    double *o = rays7 + 7 * i;
    o[0] = r.o.x, o[1] = r.o.y, o[2] = r.o.z, o[3] = r.d.x, o[4] = r.d.y, o[5] = r.d.z, o[6] = r.tm;
  }
  return 0;
}

int emu_texture_sample(void *h, int tex, const double *uvp5, size_t n, double *rgb) {
  const DevScene &S = static_cast<EmuScene *>(h)->S;
  for (size_t i = 0; i < n; i++) {
    const double *q = uvp5 + 5 * i;
    const V c = tex_sample(S, tex, q[0], q[1], v3(q[2], q[3], q[4]));
    rgb[3 * i] = c.x, rgb[3 * i + 1] = c.y, rgb[3 * i + 2] = c.z;
  }
  return 0;
}

// the body of k_scatter (shade_kernels.cuh): shade_hit on supplied records and supplied draws
int emu_scatter(void *h, const double *rays7, const double *rec12, const double *draws6, size_t n, double *out16) {
  const DevScene &S = static_cast<EmuScene *>(h)->S;
  for (size_t i = 0; i < n; i++) {
    const double *q = rays7 + 7 * i, *hr = rec12 + 12 * i, *u = draws6 + 6 * i;
    const Ray r{v3(q[0], q[1], q[2]), v3(q[3], q[4], q[5]), q[6]};
    Rec rec;
    rec.p = v3(hr[0], hr[1], hr[2]), rec.normal = v3(hr[3], hr[4], hr[5]);
    rec.t = hr[6], rec.u = hr[7], rec.v = hr[8];
    rec.front_face = hr[9] != 0.0;
    rec.mat = (int)hr[10];
    rec.prim = -1;
    FixedDraws fd;
    fd.m = Draws{u[0], u[1], u[2], 0.0};
    fd.s = Draws{u[3], u[4], u[5], 0.0};
    ScatterDetail det;
    const Scatter sc = shade_hit<-1, FixedDraws>(S, rec, r, fd, 0, &det);
    double *o = out16 + 16 * i;
    o[0] = (double)det.mode;
    o[1] = det.emitted.x, o[2] = det.emitted.y, o[3] = det.emitted.z;
    o[4] = det.attenuation.x, o[5] = det.attenuation.y, o[6] = det.attenuation.z;
    o[7] = sc.dir.x, o[8] = sc.dir.y, o[9] = sc.dir.z;
    o[10] = det.pdf_value, o[11] = det.p_scattered;
    const V w = det.mode == 0 ? v3(0, 0, 0) : sc.value;
    o[12] = w.x, o[13] = w.y, o[14] = w.z, o[15] = 0.0;
  }
  return 0;
}

// per-sample radiance, one path at a time, with the update rule of k_shade (shade_kernels.cuh)
int emu_render(void *h, int mirror, const b200rt_camera *c, uint64_t seed, int spp_begin, int spp_end, int kill_zero, double *rgb_sum,
               double *per_sample, uint64_t *rays_out) {
  const DevScene &S = static_cast<EmuScene *>(h)->S;
  const DevCamera cam = make_camera(*c);
  const size_t npix = (size_t)c->image_width * c->image_height;
  uint64_t rays = 0;
  for (int k = spp_begin; k < spp_end; k++)
    for (uint32_t pixel = 0; pixel < npix; pixel++) {
      const Stream st{seed, pixel, (uint32_t)k};
      Ray r = generate_ray(cam, st, (int)(pixel / c->image_width), (int)(pixel % c->image_width));
      V T = v3(1, 1, 1), L = v3(0, 0, 0);
      if (!(isnan(r.d.x) || isnan(r.d.y) || isnan(r.d.z))) {
        for (int bounce = 0; bounce < c->max_depth; bounce++) {
          rays++;
          Rec rec;
          bool ok;
          if (mirror) {
            MirrorCtx cx{&S, st, bounce};
            ok = mirror_hit(cx, (int)S.root, r, 0.001, B200_INF, rec, 0);
          } else {
            const Best b = trace_closest(S, r, 0.001, B200_INF, st, bounce);
            ok = b.entry >= 0;
            if (ok) finish_hit(S, S.entries[b.entry], r, b.t, rec);
          }
          if (!ok) {
            L = T * miss_color(S, r);
            break;
          }
          const Scatter sc = shade_hit(S, rec, r, st, bounce);
          if (sc.ends) {
            L = T * sc.value;
            break;
          }
          T = T * sc.value;
          if (kill_zero && T.x == 0.0 && T.y == 0.0 && T.z == 0.0) break;
          r = Ray{rec.p, sc.dir, r.tm};
        }
      }
      const double cc[3] = {L.x, L.y, L.z};
      for (int q = 0; q < 3; q++) {
        if (rgb_sum) rgb_sum[3 * (size_t)pixel + q] += cc[q];
        if (per_sample) per_sample[((size_t)(k - spp_begin) * npix + pixel) * 3 + q] = cc[q];
      }
    }
  if (rays_out) *rays_out = rays;
  return 0;
}

} // extern "C"
