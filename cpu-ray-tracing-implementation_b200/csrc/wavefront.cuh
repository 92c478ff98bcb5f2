// wavefront.cuh — the sample loop of camera::render (src/camera.h:158-172) as a wavefront of paths.
//
// A pool of `capacity` path slots lives in HBM as structure-of-arrays (coalesced 8-byte columns).  One wave is
//   k_generate  : refill free slots with fresh camera rays (generate_ray per camera mode, camera.cuh)
//   k_extend    : closest hit of every live ray (traverse.cuh), then sort the ray's index into the queue of
//                 the material kind it hit (or the miss queue) with warp ballot/popc aggregated appends
//   k_shade<K>  : one launch per material kind over its queue: rebuild the hit record, scatter / emit
//                 (shading.cuh), accumulate finished paths into the per-pixel fp64 sums and append the
//                 survivors, compacted, to the other pool
//   k_ctl       : one thread rolls the counters over
// No host round trip inside a wave: grids are fixed and every kernel reads its trip count from WaveCtl.
#pragma once
#include "camera.cuh"
#include "mirror.cuh"
#include "shading.cuh"
#include "traverse.cuh"

namespace b200 {

#define B200_NQUEUES 7 /* six material kinds (b200rt_material_kind) + miss */
#define B200_Q_MISS 6
#define B200_DEAD 0xFFFFFFFFu

struct WaveCtl {
  unsigned long long next_sample, total_samples, rays, nan_samples, waves, extend_rays_dummy;
  uint32_t n_cur, n_next, capacity, fetch; // fetch: next unclaimed ray of the persistent extend kernel
  uint32_t q_count[8];
};

struct PathBuf { // one pool: 13 columns of `capacity` elements
  double *ox, *oy, *oz, *dx, *dy, *dz, *tm, *tx, *ty, *tz;
  uint32_t *pixel, *sample, *bounce;
};

struct WaveParams {
  uint64_t seed;
  uint32_t npix, width;
  int32_t spp_begin, max_depth, kill_zero, refill; // refill: idle lanes per warp that trigger a fetch in k_extend_persistent
};

DEV Ray load_ray(const PathBuf &b, uint32_t i) { return Ray{v3(b.ox[i], b.oy[i], b.oz[i]), v3(b.dx[i], b.dy[i], b.dz[i]), b.tm[i]}; }
DEV void store_ray(const PathBuf &b, uint32_t i, const Ray &r) {
  b.ox[i] = r.o.x, b.oy[i] = r.o.y, b.oz[i] = r.o.z;
  b.dx[i] = r.d.x, b.dy[i] = r.d.y, b.dz[i] = r.d.z;
  b.tm[i] = r.tm;
}

// warp-aggregated append: one atomic per warp per queue; must be reached by all 32 lanes
DEV uint32_t warp_append(uint32_t *counter, bool want) {
  const unsigned mask = __ballot_sync(0xffffffffu, want); // callers keep whole warps in the loop
  if (!want) return 0;
  const int lane = threadIdx.x & 31;
  const int leader = __ffs(mask) - 1;
  uint32_t base = 0;
  if (lane == leader) base = atomicAdd(counter, (uint32_t)__popc(mask));
  base = __shfl_sync(mask, base, leader);
  return base + (uint32_t)__popc(mask & ((1u << lane) - 1u));
}

// ------------------------------------------------------------------------------------------------ ctl
__global__ void k_ctl_init(WaveCtl *c, unsigned long long total, uint32_t capacity) {
  c->next_sample = 0;
  c->total_samples = total;
  c->rays = 0;
  c->nan_samples = 0;
  c->waves = 0;
  c->n_cur = 0;
  c->n_next = 0;
  c->capacity = capacity;
  c->fetch = 0;
  for (int k = 0; k < 8; k++) c->q_count[k] = 0;
}
DEV uint32_t wave_refill(const WaveCtl *c) {
  const unsigned long long room = c->capacity - c->n_cur, left = c->total_samples - c->next_sample;
  return (uint32_t)(room < left ? room : left);
}
__global__ void k_ctl_generated(WaveCtl *c) {
  const uint32_t n = wave_refill(c);
  c->n_cur += n;
  c->next_sample += n;
  c->fetch = 0;
}
__global__ void k_ctl_swap(WaveCtl *c) {
  unsigned long long r = 0;
  for (int k = 0; k < B200_NQUEUES; k++) r += c->q_count[k], c->q_count[k] = 0;
  c->rays += r;
  c->n_cur = c->n_next;
  c->n_next = 0;
  c->waves++;
}

// ------------------------------------------------------------------------------------------------ generate
__global__ void __launch_bounds__(256) k_generate(const WaveCtl *c, PathBuf cur, DevCamera cam, WaveParams wp, unsigned long long *nan_samples) {
  const uint32_t n_new = wave_refill(c), n_cur = c->n_cur;
  const unsigned long long g0 = c->next_sample;
  for (uint32_t idx = blockIdx.x * blockDim.x + threadIdx.x; idx < n_new; idx += gridDim.x * blockDim.x) {
    const unsigned long long g = g0 + idx;
    const uint32_t pixel = (uint32_t)(g % wp.npix);
    const uint32_t sample = (uint32_t)wp.spp_begin + (uint32_t)(g / wp.npix);
    const Stream st{wp.seed, pixel, sample};
    const Ray r = generate_ray(cam, st, (int)(pixel / wp.width), (int)(pixel % wp.width));
    const uint32_t slot = n_cur + idx;
    const bool bad = isnan(r.d.x) || isnan(r.d.y) || isnan(r.d.z); // fisheye outside the image circle: contributes exactly 0
    store_ray(cur, slot, r);
    cur.tx[slot] = 1.0, cur.ty[slot] = 1.0, cur.tz[slot] = 1.0;
    cur.pixel[slot] = bad ? B200_DEAD : pixel;
    cur.sample[slot] = sample;
    cur.bounce[slot] = 0;
    if (bad) atomicAdd(nan_samples, 1ull);
  }
}

// ------------------------------------------------------------------------------------------------ extend
DEV int entry_material(const DevScene &S, const Entry &e) {
  switch (entry_kind(e)) {
  case E_SPHERE:
  case E_SPHERE_MOVING: return S.sphere_mat[e.row];
  case E_QUAD: return S.quad_mat[e.row];
  case E_TRI: return S.tri_mat[e.row];
  default: return S.volumes[S.vol_inst[e.row].volume].material;
  }
}

struct Queues {
  uint32_t *q[B200_NQUEUES];
};

DEV void classify(WaveCtl *c, const Queues &Q, uint32_t i, int which) {
#pragma unroll
  for (int k = 0; k < B200_NQUEUES; k++) {
    const bool mine = which == k;
    const uint32_t pos = warp_append(&c->q_count[k], mine);
    if (mine) Q.q[k][pos] = i;
  }
}

template <bool MIRROR>
__global__ void __launch_bounds__(128) k_extend(WaveCtl *c, DevScene S, PathBuf cur, HitRef *hits, Rec *recs, Queues Q, WaveParams wp) {
  const uint32_t n = c->n_cur;
  // whole warps iterate together so that the ballots in classify() see a full mask
  for (uint32_t base = (blockIdx.x * blockDim.x + threadIdx.x) & ~31u; base < n; base += gridDim.x * blockDim.x) {
    const uint32_t i = base + (threadIdx.x & 31);
    int which = -1;
    if (i < n) {
      const uint32_t pixel = cur.pixel[i];
      if (pixel != B200_DEAD) {
        const Ray r = load_ray(cur, i);
        const Stream st{wp.seed, pixel, cur.sample[i]};
        const int bounce = (int)cur.bounce[i];
        HitRef h;
        if (MIRROR) {
          MirrorCtx cx{&S, st, bounce};
          Rec rec;
          const bool ok = mirror_hit(cx, (int)S.root, r, 0.001, B200_INF, rec, 0);
          h.t = ok ? rec.t : 0.0;
          h.entry = ok ? 0 : -1;
          if (ok) recs[i] = rec;
          which = ok ? S.materials[rec.mat].kind : B200_Q_MISS;
        } else {
          const Best b = trace_closest(S, r, 0.001, B200_INF, st, bounce);
          h.t = b.t;
          h.entry = b.entry;
          which = b.entry < 0 ? B200_Q_MISS : S.materials[entry_material(S, S.entries[b.entry])].kind;
        }
        hits[i] = h;
      }
    }
    classify(c, Q, i, which);
  }
}


// The compacted-BVH extend as a persistent kernel (one CTA set sized to the SM count, no grid-stride loop):
// every lane owns one ray; a warp whose idle lanes reach `wp.refill` claims that many new rays with ONE atomic and
// continues, so lanes do not wait for the longest traversal of a fixed 32-ray batch (incoherent bounce rays differ
// by an order of magnitude in node visits).  Inner structure: "while nodes / then one leaf" so that the binary32 box
// filter and the binary64 leaf tests each run with as many lanes as possible.
__global__ void __launch_bounds__(128, 4) k_extend_persistent(WaveCtl *c, DevScene S, PathBuf cur, HitRef *hits, Queues Q, WaveParams wp) {
  const uint32_t n = c->n_cur;
  const int lane = threadIdx.x & 31;
  const unsigned FULL = 0xffffffffu;
  const int refill = wp.refill, keep = 33 - wp.refill; // leave the traversal loop when fewer than `keep` lanes still run
  bool active = false, pending = false, exhausted = false;
  uint32_t idx = 0;
  Ray r;
  RaySetup rs;
  Best best;
  Stream st;
  int bounce = 0, sp = 0;
  float lo32 = 0, hi32 = 0;
  int32_t node = B200_DONE;
  int32_t stack[B200_STACK];
  for (;;) {
    __syncwarp();
    if (__any_sync(FULL, pending)) { // retire finished rays: hit reference + index into the queue of what they hit
      int which = -1;
      if (pending) {
        HitRef h;
        h.t = best.entry < 0 ? 0.0 : best.t;
        h.entry = best.entry;
        hits[idx] = h;
        which = best.entry < 0 ? B200_Q_MISS : S.materials[entry_material(S, S.entries[best.entry])].kind;
        pending = false;
      }
      classify(c, Q, idx, which);
    }
    const unsigned idle = __ballot_sync(FULL, !active);
    if (!exhausted && __popc(idle) >= refill) {
      const uint32_t want = (uint32_t)__popc(idle);
      uint32_t base = 0;
      if (lane == 0) base = atomicAdd(&c->fetch, want);
      base = __shfl_sync(FULL, base, 0);
      if (!active) {
        const uint32_t i = base + (uint32_t)__popc(idle & ((1u << lane) - 1u));
        if (i < n) {
          const uint32_t pixel = cur.pixel[i];
          if (pixel != B200_DEAD) {
            idx = i;
            r = load_ray(cur, i);
            st = Stream{wp.seed, pixel, cur.sample[i]};
            bounce = (int)cur.bounce[i];
            rs = make_setup(S, r);
            lo32 = __double2float_rd(0.001 * rs.tsc);
            hi32 = __int_as_float(0x7f800000);
            best.t = B200_INF, best.entry = -1, best.order = 0;
            sp = 0;
            node = S.bvh_root;
            active = true;
          }
        }
      }
      exhausted = base + want >= n;
    }
    if (!__any_sync(FULL, active)) {
      if (exhausted) break;
      continue;
    }
    while (active) {
      while (node >= 0) {
        node = node_step(S, node, rs, lo32, hi32, stack, sp);
        if (node == B200_DONE && sp > 0) node = stack[--sp];
      }
      if (node != B200_DONE) {
        leaf_test(S, node, r, 0.001, B200_INF, st, bounce, best);
        if (best.entry >= 0) hi32 = __double2float_ru(best.t * rs.tsc);
        node = sp > 0 ? stack[--sp] : B200_DONE;
      }
      if (node == B200_DONE) {
        active = false;
        pending = true;
      } else if (!exhausted && __popc(__activemask()) < keep) {
        break;
      }
    }
  }
}

} // namespace b200
