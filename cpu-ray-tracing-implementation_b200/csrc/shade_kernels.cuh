// shade_kernels.cuh — material-sorted shading, accumulation, and the batch entry points
// (b200rt_trace / b200rt_generate_rays / b200rt_finalize kernels).
#pragma once
#include "wavefront.cuh"

namespace b200 {

// The shade kernels are chains of dependent gathers (queue index -> path columns + hit reference -> entry -> primitive ->
// material -> texture) at 32 warps per SM.  Experiment (-DB200_SHADE_PREFETCH=1): while a CTA works on one 128-path slice,
// every thread asks L2 (prefetch.global.L2) for the seven sectors its NEXT slice will gather.  Measured on B200: 5 % SLOWER
// on RTOW (5.34 vs 5.60 Grays/s, profiles/r01_k_layout.txt) - the kernels are limited by memory requests in flight, not by
// the latency of each, and the prefetches are extra requests.  Off.
#ifndef B200_SHADE_PREFETCH
#define B200_SHADE_PREFETCH 0
#endif
DEV void prefetch_l2(const void *p) {
#ifndef B200_HOST_EMU
  asm volatile("prefetch.global.L2 [%0];" ::"l"(p));
#endif
}
// Experiment (-DB200_OCTANT_SORT_EXPERIMENT=1, then B200RT_OPT_OCTANT_SORT): survivors of a CTA iteration written grouped by the
// octant of their new direction.  Measured (profiles/r02_j_octant_grouping.jsonl): within +-1 % on every scene, and the extra
// code costs the 64-register shade kernels spills (+5 % on the lambertian kernel, profiles/r02_t_*): compiled out.
#ifndef B200_OCTANT_SORT_EXPERIMENT
#define B200_OCTANT_SORT_EXPERIMENT 0
#endif
#ifndef B200_SHADE_MINB
#define B200_SHADE_MINB 8
#endif
// One launch per queue.  KIND = b200rt_material_kind of every hit in the queue, or B200_Q_MISS.
template <int KIND, bool MIRROR, bool EXT>
__global__ void __launch_bounds__(128, B200_SHADE_MINB) k_shade(WaveCtl *c, DevScene S, PathBuf cur, PathBuf nxt, const HitRef *hits, const Rec *recs,
                                                const uint32_t *queue, double *accum, WaveParams wp) {
  // survivors are appended to the other pool with ONE global atomic per CTA iteration (same-address atomics are the
  // scarce resource of this kernel family: see k_classify); double-buffered by iteration parity
  __shared__ uint32_t s_cnt[2][4], s_base[2];
#if B200_OCTANT_SORT_EXPERIMENT
  __shared__ uint32_t s_oct[2][8][4];
#endif
  const uint32_t n = c->q_count[KIND];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  int par = 0;
  // the queue index of the NEXT slice is fetched one iteration ahead (one register), which takes the first of the dependent
  // loads (queue -> path -> hit -> entry -> ...) off every iteration's critical path
  uint32_t i_next = blockIdx.x * 128 + threadIdx.x < n ? queue[blockIdx.x * 128 + threadIdx.x] : 0u;
  for (uint32_t bbase = blockIdx.x * 128; bbase < n; bbase += gridDim.x * 128, par ^= 1) {
    const uint32_t j = bbase + threadIdx.x;
    const uint32_t i_cur = i_next;
    {
      const uint32_t jn = j + gridDim.x * 128;
      if (jn < n) i_next = queue[jn];
    }
#if B200_SHADE_PREFETCH
    {
      const uint32_t jn = j + gridDim.x * 128;
      if (jn < n) {
        const uint32_t in = queue[jn];
        prefetch_l2(quad_ptr(cur, 0, in)), prefetch_l2(quad_ptr(cur, 1, in)), prefetch_l2(quad_ptr(cur, 2, in)), prefetch_l2(quad_ptr(cur, 3, in));
        prefetch_l2(quad_ptr(cur, 6, in)), prefetch_l2(quad_ptr(cur, 7, in));
        if (KIND != B200_Q_MISS && !MIRROR) prefetch_l2(hits + in);
      }
    }
#endif
    bool survives = false;
    Ray out_ray;
    V T;
    uint32_t pixel = 0, sample = 0, bounce = 0;
    if (j < n) {
      const uint32_t i = i_cur;
      const PathHead ph = load_head(cur, i, &T);
      const Ray r = ph.r;
      pixel = ph.pixel;
      sample = ph.sample;
      bounce = ph.bounce;
      if (KIND == B200_Q_MISS) {
        accumulate(accum, pixel, T * miss_color<EXT>(S, r));
      } else {
        Rec rec;
        if (MIRROR) {
          rec = recs[i];
        } else {
          const HitRef h = hits[i];
          finish_hit<EXT>(S, S.entries[h.entry], r, h.t, rec, B200_UV_IF_USED);
        }
        const Stream st{wp.seed, pixel, sample};
        const Scatter sc = shade_hit<KIND, Stream, EXT>(S, rec, r, st, (int)bounce);
        if (sc.ends) {
          accumulate(accum, pixel, T * sc.value);
        } else {
          T = T * sc.value;
          bounce += 1;
          // ray_color(depth <= 0) returns 0 before tracing (src/camera.h:194-195); a path whose throughput
          // is exactly 0 can only add 0 (optional cut, B200RT_OPT_KILL_ZERO_WEIGHT)
          survives = (int)bounce < wp.max_depth && !(wp.kill_zero && T.x == 0.0 && T.y == 0.0 && T.z == 0.0);
          out_ray = Ray{rec.p, sc.dir, r.tm};
        }
      }
    }
    const unsigned smask = __ballot_sync(0xffffffffu, survives);
    uint32_t pos;
#if B200_OCTANT_SORT_EXPERIMENT
    if (wp.octant_sort) {
      // The survivors of this CTA iteration (<= 128 neighbouring paths) are written grouped by the OCTANT of their new direction,
      // so that the 32-ray batches the extend kernel hands to a warp walk the tree in a similar order.  No extra memory traffic:
      // only the slot a path gets inside the CTA's block of the next pool changes.
      const int oct = survives ? ((out_ray.d.x < 0 ? 1 : 0) | (out_ray.d.y < 0 ? 2 : 0) | (out_ray.d.z < 0 ? 4 : 0)) : 8;
      uint32_t before = 0; // survivors of my warp in my octant on lower lanes
#pragma unroll
      for (int o = 0; o < 8; o++) {
        const unsigned m = __ballot_sync(0xffffffffu, oct == o);
        if (lane == 0) s_oct[par][o][warp] = (uint32_t)__popc(m);
        if (oct == o) before = (uint32_t)__popc(m & ((1u << lane) - 1u));
      }
      __syncthreads();
      if (warp == 0) { // exclusive scan over (octant major, warp minor): 32 counters, one per lane
        const uint32_t v = s_oct[par][lane >> 2][lane & 3];
        uint32_t incl = v;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
          const uint32_t t = __shfl_up_sync(0xffffffffu, incl, d);
          if (lane >= d) incl += t;
        }
        s_oct[par][lane >> 2][lane & 3] = incl - v;
        if (lane == 31) s_base[par] = incl ? atomicAdd(&c->n_next, incl) : 0;
      }
      __syncthreads();
      pos = s_base[par] + (oct < 8 ? s_oct[par][oct][warp] : 0u) + before;
    } else
#endif
    {
      if (lane == 0) s_cnt[par][warp] = (uint32_t)__popc(smask);
      __syncthreads();
      if (threadIdx.x == 0) {
        const uint32_t tot = s_cnt[par][0] + s_cnt[par][1] + s_cnt[par][2] + s_cnt[par][3];
        s_base[par] = tot ? atomicAdd(&c->n_next, tot) : 0;
      }
      __syncthreads();
      pos = s_base[par] + (uint32_t)__popc(smask & ((1u << lane) - 1u));
      for (int w = 0; w < warp; w++) pos += s_cnt[par][w];
    }
    if (survives) {
      store_path(nxt, pos, out_ray, T, pixel, sample, bounce, make_setup(wp.bound, out_ray));
    }
  }
}

// ------------------------------------------------------------------------------------------------ tail
// The last paths of a frame: once every sample has been generated and at most `limit` paths are alive, each remaining path
// is followed to its end by one thread (trace -> shade -> trace ...) instead of ten launches per nearly empty wave (with
// max_depth 50 most waves of a frame carry a fraction of a percent of the paths).  Same functions, same Philox draws
// (keyed by pixel, sample, bounce), so the contribution of every path is unchanged.  The last CTA to finish empties the
// wave (n_cur = 0): the kernels that follow in the stream find nothing to do.  WaveCtl::tail_done counts the finished CTAs
// (zeroed by k_ctl_init / k_ctl_generated; no other kernel touches it).
template <bool EXT>
__global__ void B200_MAXNREG(224) k_tail(WaveCtl *c, DevScene S, PathBuf cur, double *accum, WaveParams wp, uint32_t limit) {
  const uint32_t n = c->n_cur;
  if (n == 0 || n > limit || c->next_sample != c->total_samples) return;
  unsigned long long my_rays = 0;
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    V T;
    const PathHead ph = load_head(cur, i, &T);
    if (ph.pixel == B200_DEAD) continue;
    Ray r = ph.r;
    const Stream st{wp.seed, ph.pixel, ph.sample};
    int bounce = (int)ph.bounce;
    for (;;) {
      const Best b = trace_closest(S, r, 0.001, B200_INF, st, bounce);
      my_rays++;
      if (b.entry < 0) {
        accumulate(accum, ph.pixel, T * miss_color<EXT>(S, r));
        break;
      }
      const Entry e = S.entries[b.entry];
      Rec rec;
      finish_hit<EXT>(S, e, r, b.t, rec, B200_UV_IF_USED);
      Scatter sc;
      switch (S.materials[entry_material(S, e)].kind) {
      case B200RT_MAT_DIFFUSE_LIGHT: sc = shade_hit<B200RT_MAT_DIFFUSE_LIGHT, Stream, EXT>(S, rec, r, st, bounce); break;
      case B200RT_MAT_LAMBERTIAN: sc = shade_hit<B200RT_MAT_LAMBERTIAN, Stream, EXT>(S, rec, r, st, bounce); break;
      case B200RT_MAT_METAL: sc = shade_hit<B200RT_MAT_METAL, Stream, EXT>(S, rec, r, st, bounce); break;
      case B200RT_MAT_DIELECTRIC: sc = shade_hit<B200RT_MAT_DIELECTRIC, Stream, EXT>(S, rec, r, st, bounce); break;
      case B200RT_MAT_GLOSS: sc = shade_hit<B200RT_MAT_GLOSS, Stream, EXT>(S, rec, r, st, bounce); break;
      default: sc = shade_hit<B200RT_MAT_ISOTROPIC, Stream, EXT>(S, rec, r, st, bounce); break;
      }
      if (sc.ends) {
        accumulate(accum, ph.pixel, T * sc.value);
        break;
      }
      T = T * sc.value;
      bounce += 1;
      if (!(bounce < wp.max_depth) || (wp.kill_zero && T.x == 0.0 && T.y == 0.0 && T.z == 0.0)) break;
      r = Ray{rec.p, sc.dir, r.tm};
    }
  }
  if (my_rays) atomicAdd(&c->rays, my_rays);
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    if (atomicAdd(&c->tail_done, 1u) == gridDim.x - 1) c->n_cur = 0;
  }
}

// ------------------------------------------------------------------------------------------------ batch entry points
template <bool MIRROR>
__global__ void __launch_bounds__(128) k_trace(DevScene S, const double *rays7, size_t n, double lo, double hi, uint64_t seed, int32_t *prim,
                                                double *t, double *uv, uint8_t *front, double *p, double *normal) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    const double *q = rays7 + 7 * i;
    const Ray r{v3(q[0], q[1], q[2]), v3(q[3], q[4], q[5]), q[6]};
    const Stream st{seed, (uint32_t)i, (uint32_t)(i >> 32)};
    Rec rec;
    bool ok;
    if (MIRROR) {
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
}

// One level of ray_color below the hit test on caller-supplied hit records and caller-supplied draws (b200rt_scatter).
// rec12 = {p[3], normal[3], t, u, v, front_face, material row, 0}; draws6 = {m0 m1 m2 s0 s1 s2};
// out16 = {mode, emitted[3], attenuation[3], dir[3], pdf_value, p_scattered, weight[3], 0} with weight = what multiplies the
// incoming radiance: attenuation (kDetermined) or attenuation * p_scattered / pdf_value (kRandom).
__global__ void __launch_bounds__(128) k_scatter(DevScene S, const double *rays7, const double *rec12, const double *draws6, size_t n, double *out16) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    const double *q = rays7 + 7 * i, *h = rec12 + 12 * i, *u = draws6 + 6 * i;
    const Ray r{v3(q[0], q[1], q[2]), v3(q[3], q[4], q[5]), q[6]};
    Rec rec;
    rec.p = v3(h[0], h[1], h[2]), rec.normal = v3(h[3], h[4], h[5]);
    rec.t = h[6], rec.u = h[7], rec.v = h[8];
    rec.front_face = h[9] != 0.0;
    rec.mat = (int)h[10];
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
}

__global__ void __launch_bounds__(256) k_raygen(DevCamera cam, uint64_t seed, int spp_begin, size_t n, double *rays7) {
  const size_t npix = (size_t)cam.c.image_width * cam.c.image_height;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    const uint32_t pixel = (uint32_t)(i % npix);
    const uint32_t sample = (uint32_t)spp_begin + (uint32_t)(i / npix);
    const Ray r = generate_ray(cam, Stream{seed, pixel, sample}, (int)(pixel / cam.c.image_width), (int)(pixel % cam.c.image_width));
    double *o = rays7 + 7 * i;
    o[0] = r.o.x, o[1] = r.o.y, o[2] = r.o.z, o[3] = r.d.x, o[4] = r.d.y, o[5] = r.d.z, o[6] = r.tm;
  }
}

// pixel_color / spp and write_color (src/camera.h:169-170, src/color.h:16-36): gamma 2.2, int(255.999 c), unclamped
__global__ void __launch_bounds__(256) k_finalize(const double *sum, size_t n, double spp, double *linear, int32_t *ppm) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    const double c = sum[i] / spp;
    if (linear) linear[i] = c;
    if (ppm) {
      const double g = c > 0 ? pow(c, 1 / 2.2) : 0.0;
      const double v = 255.999 * g;
      ppm[i] = (v > -2147483649.0 && v < 2147483648.0) ? (int32_t)v : (-2147483647 - 1); // NaN / out of range: what x86 cvttsd2si gives the reference
    }
  }
}

// texture::sample on a batch (parity entry for tests of the texture path)
__global__ void __launch_bounds__(128) k_texture(DevScene S, int tex, const double *uvp5, size_t n, double *rgb) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    const double *q = uvp5 + 5 * i;
    const V c = tex_sample(S, tex, q[0], q[1], v3(q[2], q[3], q[4]));
    rgb[3 * i] = c.x, rgb[3 * i + 1] = c.y, rgb[3 * i + 2] = c.z;
  }
}

} // namespace b200
