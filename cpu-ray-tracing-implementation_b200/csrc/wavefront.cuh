// wavefront.cuh — the sample loop of camera::render (src/camera.h:158-172) as a wavefront of paths.
//
// A pool of `capacity` path slots lives in HBM as structure-of-arrays (four columns of 32 bytes per path).  One wave is
//   k_generate  : refill free slots with fresh camera rays (generate_ray per camera mode, camera.cuh)
//   k_extend    : closest hit of every live ray (traverse.cuh): persistent warps, rays staged through shared memory
//   k_classify  : sort every path's index into the queue of the material kind it hit (or the miss queue):
//                 ballot/popc ranking per warp, one global atomic per queue per 2048-path chunk
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
#define B200_Q_MISS_DONE 7 /* a miss that the extend kernel has already accounted for (WaveParams::fused_miss): counted, never queued */
#define B200_DEAD 0xFFFFFFFFu

struct WaveCtl {
  unsigned long long next_sample, total_samples, rays, nan_samples, waves, wave_live; // wave_live: this wave had paths to trace (an empty trailing wave of the graph loop is not counted)
  uint32_t n_cur, n_next, capacity, fetch; // fetch: next unclaimed 32-ray batch of the persistent extend kernel
  uint32_t q_count[8];
  uint32_t n_gen, tail_done; // n_gen: live paths appended by k_generate<COMPACT> in this wave; tail_done: CTAs of k_tail that have finished
};

// Path state = eight 16-byte "quads" per path.  The fp32 box-filter constants (traverse.cuh) are computed by the PRODUCER
// of the ray (k_generate / k_shade) and ride along in quads 4-6, so the extend kernel starts traversing straight away.
//   q0 {o.x o.y}  q1 {o.z d.x}  q2 {d.y d.z}  q3 {time T.x}  q4 {ix iy iz c0x}  q5 {c1x c0y c1y c0z}
//   q6 {c1z pixel sample bounce}  q7 {T.y T.z}
#define B200_PATH_QUADS 8
struct PathBuf {
  double2 *base;
  uint32_t capacity;
};
// B200_PATH_PAIRED=0: quad q of path i at q * capacity + i (eight columns of 16 bytes).
// B200_PATH_PAIRED=1: quads 2k and 2k+1 of one path share a 32-byte sector (four columns of 32 bytes) and are moved by ONE
// 256-bit access (LDG/STG.E.ENL2.256 on sm_100a): the queue-ordered gathers of the shade kernels read q0..q3, q6, q7 as three
// full sectors with three requests instead of six half-used sectors with six; producers write four 32-byte stores per path.
// (Pairing WITHOUT the 256-bit accesses was measured 4.5 % slower than columns: 16 bytes per lane at a 32-byte stride.)
#ifndef B200_PATH_PAIRED
#define B200_PATH_PAIRED 1
#endif
DEV size_t quad_index(uint32_t capacity, int q, uint32_t i) {
#if B200_PATH_PAIRED
  return ((size_t)(q >> 1) * capacity + i) * 2 + (size_t)(q & 1);
#else
  return (size_t)q * capacity + i;
#endif
}
DEV const double2 *quad_ptr(const PathBuf &b, int q, uint32_t i) { return b.base + quad_index(b.capacity, q, i); }
// The two pools of a pipeline: wave k reads pool k & 1 and appends its survivors to the other one.  The kernels get the pools
// of their wave as plain by-value parameters (constant bank: no registers).  Selecting the pool on the device from the wave
// counter was tried for the CUDA-graph loop and cost the 64-register shade kernels 40-60 bytes of extra spills (+5 % on the
// lambertian kernel, profiles/r02_t_*); the graph's loop body holds an even and an odd wave instead (api.cu).
struct Pools {
  PathBuf b[2];
};

struct WaveParams {
  uint64_t seed;
  uint32_t npix, width;
  int32_t spp_begin, max_depth, kill_zero, refill; // refill: idle lanes per warp that trigger a fetch in k_extend_persistent
  int32_t octant_sort; // k_shade: survivors of a CTA iteration are written grouped by the octant of their new direction
  int32_t leaf_lanes; // k_extend_persistent<UNI>: lanes waiting at a leaf that trigger the leaf phase (32 = only when no lane is at a node)
  int32_t fused_miss; // 1 = the scene has no background: a miss adds 0, the extend kernel finishes it itself (no miss queue, no miss launch)
  double bound[3]; // DevScene::bound, for the producers' make_setup
};

struct PathHead { // what every consumer needs
  Ray r;
  uint32_t pixel, sample, bounce;
};
#if B200_PATH_PAIRED
// one 256-bit access = two adjacent quads; p must be 32-byte aligned
// B200_STREAM_HINTS=1: the path state is written once and read once per wave and a wave (gigabytes) is far larger than the
// 126 MB L2, so its lines only push the scene (BVH, primitives: re-read by every ray) out of L2.  They are marked
// evict-first in L2 and not allocated in L1 (LDG/STG.E.NA.EFL2.256, LDGSTS with an evict-first policy).
#ifndef B200_STREAM_HINTS
#define B200_STREAM_HINTS 1
#endif
DEV void ld256(const double2 *p, double2 &a, double2 &b) {
#if B200_STREAM_HINTS
  asm volatile("ld.global.nc.L1::no_allocate.L2::evict_first.v4.f64 {%0,%1,%2,%3}, [%4];" : "=d"(a.x), "=d"(a.y), "=d"(b.x), "=d"(b.y) : "l"(p));
#else
  asm volatile("ld.global.nc.v4.f64 {%0,%1,%2,%3}, [%4];" : "=d"(a.x), "=d"(a.y), "=d"(b.x), "=d"(b.y) : "l"(p));
#endif
}
DEV void st256(double2 *p, const double2 &a, const double2 &b) {
#if B200_STREAM_HINTS
  asm volatile("st.global.L1::no_allocate.L2::evict_first.v4.f64 [%0], {%1,%2,%3,%4};" ::"l"(p), "d"(a.x), "d"(a.y), "d"(b.x), "d"(b.y) : "memory");
#else
  asm volatile("st.global.v4.f64 [%0], {%1,%2,%3,%4};" ::"l"(p), "d"(a.x), "d"(a.y), "d"(b.x), "d"(b.y) : "memory");
#endif
}
DEV double pack2(float lo, float hi) { return __hiloint2double(__float_as_int(hi), __float_as_int(lo)); }
DEV float unpack_lo(double d) { return __int_as_float(__double2loint(d)); }
DEV float unpack_hi(double d) { return __int_as_float(__double2hiint(d)); }
#endif
// ray, ids and throughput of path i: everything a shade kernel gathers
DEV PathHead load_head(const PathBuf &b, uint32_t i, V *T = nullptr) {
  PathHead h;
#if B200_PATH_PAIRED
  double2 q0, q1, q2, q3, q6, q7;
  ld256(quad_ptr(b, 0, i), q0, q1);
  ld256(quad_ptr(b, 2, i), q2, q3);
  ld256(quad_ptr(b, 6, i), q6, q7);
  h.r = Ray{v3(q0.x, q0.y, q1.x), v3(q1.y, q2.x, q2.y), q3.x};
  h.pixel = (uint32_t)__double2hiint(q6.x), h.sample = (uint32_t)__double2loint(q6.y), h.bounce = (uint32_t)__double2hiint(q6.y);
  if (T) *T = v3(q3.y, q7.x, q7.y);
#else
  const double2 q0 = __ldg(quad_ptr(b, 0, i)), q1 = __ldg(quad_ptr(b, 1, i)), q2 = __ldg(quad_ptr(b, 2, i)), q3 = __ldg(quad_ptr(b, 3, i));
  const int4 q6 = __ldg(reinterpret_cast<const int4 *>(quad_ptr(b, 6, i)));
  h.r = Ray{v3(q0.x, q0.y, q1.x), v3(q1.y, q2.x, q2.y), q3.x};
  h.pixel = (uint32_t)q6.y, h.sample = (uint32_t)q6.z, h.bounce = (uint32_t)q6.w;
  if (T) {
    const double2 q7 = __ldg(quad_ptr(b, 7, i));
    *T = v3(q3.y, q7.x, q7.y);
  }
#endif
  return h;
}
DEV void store_path(const PathBuf &b, uint32_t i, const Ray &r, const V &T, uint32_t pixel, uint32_t sample, uint32_t bounce, const RaySetup &s) {
  double2 *q = b.base;
  const uint32_t n = b.capacity;
#if B200_PATH_PAIRED
  st256(q + quad_index(n, 0, i), make_double2(r.o.x, r.o.y), make_double2(r.o.z, r.d.x));
  st256(q + quad_index(n, 2, i), make_double2(r.d.y, r.d.z), make_double2(r.tm, T.x));
  st256(q + quad_index(n, 4, i), make_double2(pack2(s.ix, s.iy), pack2(s.iz, s.c0x)), make_double2(pack2(s.c1x, s.c0y), pack2(s.c1y, s.c0z)));
  st256(q + quad_index(n, 6, i), make_double2(pack2(s.c1z, __uint_as_float(pixel)), pack2(__uint_as_float(sample), __uint_as_float(bounce))), make_double2(T.y, T.z));
#else
  q[quad_index(n, 0, i)] = make_double2(r.o.x, r.o.y);
  q[quad_index(n, 1, i)] = make_double2(r.o.z, r.d.x);
  q[quad_index(n, 2, i)] = make_double2(r.d.y, r.d.z);
  q[quad_index(n, 3, i)] = make_double2(r.tm, T.x);
  *reinterpret_cast<float4 *>(q + quad_index(n, 4, i)) = make_float4(s.ix, s.iy, s.iz, s.c0x);
  *reinterpret_cast<float4 *>(q + quad_index(n, 5, i)) = make_float4(s.c1x, s.c0y, s.c1y, s.c0z);
  *reinterpret_cast<float4 *>(q + quad_index(n, 6, i)) = make_float4(s.c1z, __uint_as_float(pixel), __uint_as_float(sample), __uint_as_float(bounce));
  q[quad_index(n, 7, i)] = make_double2(T.y, T.z);
#endif
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
  c->wave_live = 0;
  c->n_cur = 0;
  c->n_next = 0;
  c->capacity = capacity;
  c->fetch = 0;
  c->n_gen = 0;
  c->tail_done = 0;
  for (int k = 0; k < 8; k++) c->q_count[k] = 0;
}
DEV uint32_t wave_refill(const WaveCtl *c) {
  const unsigned long long room = c->capacity - c->n_cur, left = c->total_samples - c->next_sample;
  return (uint32_t)(room < left ? room : left);
}
__global__ void k_ctl_generated(WaveCtl *c, int compact) {
  const uint32_t n = wave_refill(c);
  c->n_cur += compact ? c->n_gen : n;
  c->next_sample += n;
  c->wave_live = c->n_cur != 0 ? 1ull : c->wave_live;
  c->fetch = 0;
  c->n_gen = 0;
  c->tail_done = 0;
}
__global__ void k_ctl_swap(WaveCtl *c) {
  unsigned long long r = 0;
  for (int k = 0; k < 8; k++) r += c->q_count[k], c->q_count[k] = 0; // 7 queues + misses finished inside extend
  c->rays += r;
  c->n_cur = c->n_next;
  c->n_next = 0;
  c->waves += c->wave_live;
  c->wave_live = 0;
}

// The same roll-over as the last node of the CUDA-graph while loop (api.cu): the loop runs another wave while paths are
// alive or samples remain to be generated.
#ifndef B200_HOST_EMU
__global__ void k_ctl_swap_cond(WaveCtl *c, cudaGraphConditionalHandle more) {
  unsigned long long r = 0;
  for (int k = 0; k < 8; k++) r += c->q_count[k], c->q_count[k] = 0;
  c->rays += r;
  c->n_cur = c->n_next;
  c->n_next = 0;
  c->waves += c->wave_live;
  c->wave_live = 0;
  cudaGraphSetConditional(more, (c->n_cur != 0 || c->next_sample < c->total_samples) ? 1u : 0u);
}
#endif

// ------------------------------------------------------------------------------------------------ generate
// COMPACT (fisheye cameras): samples outside the image circle produce no ray (NaN direction, src/camera.h:268-282) and add
// exactly 0; they are counted and DROPPED here instead of riding through extend and classify as dead slots (56 % of the Fox
// config's samples).  Live paths are appended behind n_cur with one atomic per CTA iteration; k_ctl_generated adds n_gen.
template <bool COMPACT>
#ifndef B200_GENERATE_POOLS
#define B200_GENERATE_POOLS 1 // the generate kernels select their pool from a by-value pair with a run-time parity: +5.7 % on the Fox config (k_generate<COMPACT> is 45 % of that frame), +-0 elsewhere (profiles/r02_w_generate_pools.jsonl)
#endif
#if B200_GENERATE_POOLS
__global__ void __launch_bounds__(256) k_generate(WaveCtl *c, Pools P, int parity, DevCamera cam, WaveParams wp) {
  const PathBuf cur = P.b[parity];
#else
__global__ void __launch_bounds__(256) k_generate(WaveCtl *c, PathBuf cur, DevCamera cam, WaveParams wp) {
#endif
  const uint32_t n_new = wave_refill(c), n_cur = c->n_cur;
  const unsigned long long g0 = c->next_sample;
  __shared__ uint32_t warp_live[8], cta_base;
  for (uint32_t first = blockIdx.x * blockDim.x; first < n_new; first += gridDim.x * blockDim.x) {
    const uint32_t idx = first + threadIdx.x;
    const bool in = idx < n_new;
    Ray r{};
    uint32_t pixel = 0, sample = 0;
    bool bad = false;
    if (in) {
      const unsigned long long g = g0 + idx;
      pixel = (uint32_t)(g % wp.npix);
      sample = (uint32_t)wp.spp_begin + (uint32_t)(g / wp.npix);
      r = generate_ray<true>(cam, Stream{wp.seed, pixel, sample}, (int)(pixel / wp.width), (int)(pixel % wp.width));
      bad = isnan(r.d.x) || isnan(r.d.y) || isnan(r.d.z); // fisheye outside the image circle: contributes exactly 0
    }
    if (COMPACT) {
      const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
      const unsigned live = __ballot_sync(0xffffffffu, in && !bad);
      if (lane == 0) warp_live[wib] = __popc(live);
      __syncthreads();
      if (threadIdx.x == 0) {
        uint32_t tot = 0;
        for (int k = 0; k < 8; k++) tot += warp_live[k];
        const uint32_t n_in = min(blockDim.x, n_new - first);
        cta_base = tot ? atomicAdd(&c->n_gen, tot) : 0u;
        if (n_in > tot) atomicAdd(&c->nan_samples, (unsigned long long)(n_in - tot));
      }
      __syncthreads();
      uint32_t pos = cta_base + __popc(live & ((1u << lane) - 1u));
      for (int k = 0; k < wib; k++) pos += warp_live[k];
      if (in && !bad) store_path(cur, n_cur + pos, r, v3(1.0, 1.0, 1.0), pixel, sample, 0, make_setup(wp.bound, r));
      __syncthreads(); // warp_live / cta_base are rewritten by the next iteration
    } else if (in) {
      store_path(cur, n_cur + idx, r, v3(1.0, 1.0, 1.0), bad ? B200_DEAD : pixel, sample, 0, make_setup(wp.bound, r));
      if (bad) atomicAdd(&c->nan_samples, 1ull);
    }
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

__global__ void k_fill_entry_queue(DevScene S, uint8_t *out) { // DevScene::entry_queue
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < S.n_entries) out[i] = (uint8_t)S.materials[entry_material(S, S.entries[i])].kind;
}

struct Queues {
  uint32_t *q[B200_NQUEUES];
};

DEV void store_hit(HitRef *hits, uint32_t i, double t, int32_t entry, int32_t queue) { // one 128-bit store
  int4 w;
  w.x = __double2loint(t), w.y = __double2hiint(t), w.z = entry, w.w = queue;
  *reinterpret_cast<int4 *>(hits + i) = w;
}

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
      const PathHead ph = load_head(cur, i);
      if (ph.pixel != B200_DEAD) {
        const Ray r = ph.r;
        const Stream st{wp.seed, ph.pixel, ph.sample};
        const int bounce = (int)ph.bounce;
        HitRef h;
        if (MIRROR) {
          MirrorCtx cx{&S, st, bounce};
          Rec rec;
          const bool ok = mirror_hit(cx, (int)S.root, r, 0.001, B200_INF, rec, 0);
          h.t = ok ? rec.t : 0.0;
          h.entry = ok ? 0 : -1;
          h.queue = -1;
          if (ok) recs[i] = rec;
          which = ok ? S.materials[rec.mat].kind : B200_Q_MISS;
        } else {
          const Best b = trace_closest(S, r, 0.001, B200_INF, st, bounce);
          h.t = b.t;
          h.entry = b.entry;
          h.queue = -1;
          which = b.entry < 0 ? B200_Q_MISS : S.materials[entry_material(S, S.entries[b.entry])].kind;
        }
        hits[i] = h;
      }
    }
    classify(c, Q, i, which);
  }
}


// ------------------------------------------------------------------------------------------------ persistent extend
DEV void accumulate(double *accum, uint32_t pixel, V c) { // image_[i*W+j] += sample, src/camera.h:167-169
  if (c.x != 0.0) atomicAdd(accum + 3 * (size_t)pixel, c.x);
  if (c.y != 0.0) atomicAdd(accum + 3 * (size_t)pixel + 1, c.y);
  if (c.z != 0.0) atomicAdd(accum + 3 * (size_t)pixel + 2, c.z);
}
DEV void cp_async16(void *smem_dst, const void *gmem_src) {
#ifndef B200_HOST_EMU
  const unsigned d = (unsigned)__cvta_generic_to_shared(smem_dst);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(d), "l"(gmem_src));
#endif
}
DEV unsigned long long l2_evict_first_policy() {
  unsigned long long pol = 0;
#ifndef B200_HOST_EMU
  asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
#endif
  return pol;
}
DEV void cp_async16_stream(void *smem_dst, const void *gmem_src, unsigned long long pol) {
#ifndef B200_HOST_EMU
  const unsigned d = (unsigned)__cvta_generic_to_shared(smem_dst);
#if B200_STREAM_HINTS
  asm volatile("cp.async.cg.shared.global.L2::cache_hint [%0], [%1], 16, %2;" ::"r"(d), "l"(gmem_src), "l"(pol));
#else
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(d), "l"(gmem_src));
#endif
#endif
}
DEV void cp_async_commit() {
#ifndef B200_HOST_EMU
  asm volatile("cp.async.commit_group;");
#endif
}
DEV void cp_async_wait_all() {
#ifndef B200_HOST_EMU
  asm volatile("cp.async.wait_group 0;" ::: "memory");
#endif
}

#define B200_EXT_QUADS 7 /* q0..q6: ray, time, filter constants, pixel/sample/bounce; the throughput (q7) stays in HBM */

// The compacted-BVH extend as a persistent kernel: warps own batches of 32 consecutive rays (warp w takes batches
// w, w + W, ...), and every batch is brought into shared memory by cp.async one batch AHEAD of its use, so the HBM
// latency of the ray fetch overlaps the traversal of the previous batch.  Every lane owns one ray; when `wp.refill`
// lanes of a warp have finished, they take the next rays from the staged batch (a few LDS.128, no global round
// trip, no atomics) instead of idling until the slowest of 32 incoherent rays is done.  A finished lane stores its
// {t, leaf} itself; sorting into material queues is a separate coalesced pass (k_classify).
// Inner structure: "while nodes / then one leaf" so that the binary32 box filter and the binary64 leaf tests each
// run with as many lanes as possible.
// SPEC: speculative traversal — a lane that reaches its first leaf postpones it and keeps walking inner nodes until it
// meets a second leaf (or runs out), so that the node phase and the leaf phase of a warp each run with more lanes; the price
// is that the postponed leaf has not yet shrunk the interval for the nodes walked in between.
// WARPS x MINB = resident warps per SM the kernel is compiled for: registers are handed out per warp, so 2-warp CTAs give
// occupancy points (18 warps at 112 registers) that 4-warp CTAs skip.
#ifdef B200_HOST_EMU
#define B200_MAXNREG(n)
#else
#define B200_MAXNREG(n) __maxnreg__(n)
#endif
// STAGES: 2 = the next batch is always in flight (needed when lanes refill mid-batch); 1 = the single stage is refilled as
// soon as its 32 rays have been taken into registers.  Shared memory is carved out of the L1 that caches the BVH, so the
// smaller footprint (14 KB instead of 28 KB per CTA) is not free to give away: see DESIGN.md.
// WIDE: the tree is the BVH4 of scene_dev.cuh (S.bvh4) instead of the BVH2.
#ifndef B200_RAY_SORT_EXPERIMENT
#define B200_RAY_SORT_EXPERIMENT 0
#endif
#ifndef B200_EXT_DYNAMIC
#define B200_EXT_DYNAMIC 4 // largest number of batches claimed per atomic; 0 = static striding (profiles/r01_v_*)
#endif
// UNI: the traversal loop is WARP-UNIFORM ("threshold if-if"): every iteration the warp votes and then either all lanes that
// stand at an inner node take one node step, or all lanes that stand at a leaf test it.  The leaf phase is entered when no
// lane is at a node or when at least wp.leaf_lanes lanes wait at a leaf; leaf_lanes = 32 is the while-while schedule (leaves
// only when every lane has found one), small values stop lanes that have found their leaf from idling through the node
// steps of the slowest lane.  Finished lanes refill from the staged batch as before.
template <int MAXREG, bool SPEC, int WARPS, int STAGES, bool WIDE, bool UNI = false>
__global__ void B200_MAXNREG(MAXREG) k_extend_persistent(const WaveCtl *c, DevScene S, PathBuf cur, HitRef *hits, WaveParams wp, const uint32_t *order) {
  // order (may be null): a permutation of the live paths; batch b then holds paths order[32 b .. 32 b + 31] instead of 32 consecutive
  // ones (the ray-reordering experiment, B200RT_OPT_RAY_SORT, compiled in with -DB200_RAY_SORT_EXPERIMENT=1): the staging becomes a
  // gather, everything after it is unchanged.  Measured (profiles/r02_n_ray_sort_potential.jsonl): a full radix sort of the rays
  // by direction and origin buys at most 11 % of the extend time (Sponza stand-in) and costs 8-13 % on the small scenes.
#if B200_RAY_SORT_EXPERIMENT
  __shared__ uint32_t stage_idx[WARPS][STAGES][32];
#else
  order = nullptr;
#endif
#if B200_PATH_PAIRED
  __shared__ __align__(32) double2 stage_mem[WARPS][STAGES][224]; // warp x stage x {3 columns x 32 paths x 2 quads, then q6 of 32 paths}: 3.5 KB
#else
  __shared__ __align__(16) double2 stage_mem[WARPS][STAGES][B200_EXT_QUADS][32]; // warp x stage x quad x lane: 3.5 KB per warp and stage
#endif
  const uint32_t n = c->n_cur;
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  const unsigned FULL = 0xffffffffu;
  const uint32_t n_warps = gridDim.x * WARPS, n_batches = (n + 31) / 32;
  const int refill = wp.refill, keep = 33 - wp.refill;
  const unsigned long long stream_pol = l2_evict_first_policy();

  auto prefetch = [&](int stg, uint32_t batch) { // one cp.async group per call, also when there is nothing to copy
#if B200_PATH_PAIRED
    // a batch is 1 KB contiguous per 32-byte column: lane l copies 16-byte chunks l and 32 + l of each (fully coalesced)
    if (batch < n_batches) {
      const uint32_t i0 = batch * 32;
#if B200_RAY_SORT_EXPERIMENT
      if (order) { // gather: slot s of the batch is path order[i0 + s]
        const uint32_t mine = i0 + lane < n ? __ldg(order + i0 + lane) : 0u;
        stage_idx[wib][stg][lane] = mine;
#pragma unroll
        for (int k = 0; k < 2; k++) {
          const uint32_t ch = 32 * k + lane, slot = ch >> 1;
          const uint32_t pidx = __shfl_sync(FULL, mine, slot);
          if (i0 + slot < n) {
#pragma unroll
            for (int c = 0; c < 3; c++) cp_async16_stream(&stage_mem[wib][stg][64 * c + ch], cur.base + ((size_t)c * cur.capacity + pidx) * 2 + (ch & 1u), stream_pol);
          }
        }
        if (i0 + lane < n) cp_async16_stream(&stage_mem[wib][stg][192 + lane], quad_ptr(cur, 6, mine), stream_pol);
      } else
#endif
      {
#pragma unroll
      for (int c = 0; c < 3; c++)
#pragma unroll
        for (int k = 0; k < 2; k++) {
          const uint32_t ch = 32 * k + lane;
          if (i0 + (ch >> 1) < n) cp_async16_stream(&stage_mem[wib][stg][64 * c + ch], cur.base + ((size_t)c * cur.capacity + i0) * 2 + ch, stream_pol);
        }
      if (i0 + lane < n) cp_async16_stream(&stage_mem[wib][stg][192 + lane], quad_ptr(cur, 6, i0 + lane), stream_pol); // q6 only: q7 (throughput) stays in HBM
      }
    }
#else
    const uint32_t i = batch * 32 + lane;
    if (batch < n_batches && i < n) {
#pragma unroll
      for (int q = 0; q < B200_EXT_QUADS; q++) cp_async16(&stage_mem[wib][stg][q][lane], quad_ptr(cur, q, i));
    }
#endif
    cp_async_commit();
  };

  // A ray that hits nothing.  Without a background camera::miss adds exactly 0 (src/camera.h:181): the path is finished
  // here (counted, never queued) and the scene needs no miss launch at all.
  auto finish_miss = [&](uint32_t i) { store_hit(hits, i, 0.0, -1, wp.fused_miss ? B200_Q_MISS_DONE : B200_Q_MISS); };

#if B200_EXT_DYNAMIC
  // batches are CLAIMED (B200_EXT_DYNAMIC at a time, one atomic per claim) instead of strided: a warp that drew cheap rays
  // takes more of them, which shortens the tail of the launch.  The atomic for the NEXT chunk is issued when the last batch
  // of the current one is handed out, so its round trip is hidden behind a whole batch of traversal.
  uint32_t cl_next = 0, cl_end = 0, cl_pending = 0;
  uint32_t *const fetch = const_cast<uint32_t *>(&c->fetch);
  // guided chunks: B200_EXT_DYNAMIC batches per claim while plenty remain (few atomics), shrinking to 1 over the last
  // 2 * B200_EXT_DYNAMIC batches per warp (fine-grained tail).  `seen` may be stale: any chunk size is valid.
  const uint32_t inv = 0xffffffffu / (2u * n_warps);
  auto chunk_for = [&](uint32_t seen) -> uint32_t {
    const uint32_t rem = seen < n_batches ? n_batches - seen : 0u;
    return min((uint32_t)B200_EXT_DYNAMIC, max(1u, __umulhi(rem, inv)));
  };
  uint32_t cl_ch = chunk_for(0);
  if (lane == 0) cl_pending = atomicAdd(fetch, cl_ch);
  auto claim = [&]() -> uint32_t {
    if (cl_next == cl_end) {
      cl_next = __shfl_sync(FULL, cl_pending, 0);
      cl_end = cl_next + cl_ch;
    }
    const uint32_t b = cl_next++;
    if (cl_next == cl_end) {
      cl_ch = chunk_for(cl_end);
      if (lane == 0) cl_pending = atomicAdd(fetch, cl_ch);
    }
    return b;
  };
  uint32_t batch = claim(), batch_ahead = 0;
#else
  uint32_t batch = blockIdx.x * WARPS + wib; // the batch staged in `stg`
#endif
  int stg = 0;
  prefetch(0, batch);
  cp_async_wait_all();
  __syncwarp();
  uint32_t base = batch * 32, count = batch < n_batches ? min(32u, n - base) : 0u, taken = 0;
#if B200_EXT_DYNAMIC
  if (STAGES == 2) prefetch(1, batch_ahead = claim());
#else
  if (STAGES == 2) prefetch(1, batch + n_warps);
#endif

  bool active = false;
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
    unsigned idle = __ballot_sync(FULL, !active);
    if (count > 0 && (__popc(idle) >= refill || idle == FULL)) {
      while (idle != 0 && count > 0) {
        if (taken == count) { // next staged batch (already in flight)
          const int drained = stg;
#if B200_EXT_DYNAMIC
          batch = batch_ahead; // STAGES == 1: set when the prefetch was issued
#else
          batch += n_warps;
#endif
          if (STAGES == 2) stg ^= 1;
          cp_async_wait_all();
          __syncwarp();
          base = batch * 32, count = batch < n_batches ? min(32u, n - base) : 0u, taken = 0;
#if B200_EXT_DYNAMIC
          if (STAGES == 2) prefetch(drained, batch_ahead = claim());
#else
          if (STAGES == 2) prefetch(drained, batch + n_warps); // refill the stage just drained
#endif
          continue;
        }
        const uint32_t slot = taken + (uint32_t)__popc(idle & ((1u << lane) - 1u));
        const bool take = !active && slot < count;
        if (take) {
#if B200_PATH_PAIRED
          const double2 *sm = &stage_mem[wib][stg][2 * slot];
          const double2 q0 = sm[0], q1 = sm[1], q2 = sm[64], q3 = sm[65];
          const float4 q4 = *reinterpret_cast<const float4 *>(sm + 128), q5 = *reinterpret_cast<const float4 *>(sm + 129);
          const float4 q6 = *reinterpret_cast<const float4 *>(&stage_mem[wib][stg][192 + slot]);
#else
          const double2 *sm = &stage_mem[wib][stg][0][slot];
          const double2 q0 = sm[0], q1 = sm[32], q2 = sm[64], q3 = sm[96];
          const float4 q4 = *reinterpret_cast<const float4 *>(sm + 128), q5 = *reinterpret_cast<const float4 *>(sm + 160);
          const float4 q6 = *reinterpret_cast<const float4 *>(sm + 192);
#endif
#if B200_RAY_SORT_EXPERIMENT
          idx = order ? stage_idx[wib][stg][slot] : base + slot;
#else
          idx = base + slot;
#endif
          const uint32_t pixel = __float_as_uint(q6.y);
          if (pixel == B200_DEAD) { // a camera sample that produced no ray: nothing to trace, nothing to shade
            store_hit(hits, idx, 0.0, -2, -1);
          } else {
            r = Ray{v3(q0.x, q0.y, q1.x), v3(q1.y, q2.x, q2.y), q3.x};
            st = Stream{wp.seed, pixel, __float_as_uint(q6.z)};
            bounce = (int)__float_as_uint(q6.w);
            rs.ix = q4.x, rs.iy = q4.y, rs.iz = q4.z, rs.c0x = q4.w;
            rs.c1x = q5.x, rs.c0y = q5.y, rs.c1y = q5.z, rs.c0z = q5.w;
            rs.c1z = q6.x;
            rs.tsc = ray_tsc(r.d);
            lo32 = __double2float_rd(0.001 * rs.tsc);
            hi32 = __int_as_float(0x7f800000);
            best.t = B200_INF, best.entry = -1, best.order = 0;
            sp = 0;
            node = S.bvh_root;
            active = node != B200_EMPTY_ROOT;
            if (!active) finish_miss(idx);
          }
        }
        taken += min((uint32_t)__popc(idle), count - taken);
        if (STAGES == 1 && taken == count) { // every ray of the stage is in registers now: start fetching the next batch into it
          __syncwarp();
#if B200_EXT_DYNAMIC
          prefetch(0, batch_ahead = claim());
#else
          prefetch(0, batch + n_warps);
#endif
        }
        idle = __ballot_sync(FULL, !active);
      }
    }
    if (!__any_sync(FULL, active)) {
      if (count == 0) break;
      continue;
    }
    if (UNI) {
      const int T = wp.leaf_lanes;
      for (;;) {
        const bool at_node = active && node >= 0, at_leaf = active && node < 0; // active lanes never hold B200_DONE here
        const unsigned mn = __ballot_sync(FULL, at_node), ml = __ballot_sync(FULL, at_leaf);
        if ((mn | ml) == 0u) break;
        if (count > 0 && 32 - __popc(mn | ml) >= refill) break; // enough finished lanes: take new rays first
        if (mn != 0u && __popc(ml) < T) {
          if (at_node) {
            node = WIDE ? node_step4(S, node, rs, lo32, hi32, stack, sp) : node_step(S, node, rs, lo32, hi32, stack, sp);
            if (node == B200_DONE && sp > 0) node = stack[--sp];
          }
        } else if (at_leaf) {
          leaf_test(S, node, r, 0.001, B200_INF, st, bounce, best);
          if (best.entry >= 0) hi32 = __double2float_ru(best.t * rs.tsc);
          node = sp > 0 ? stack[--sp] : B200_DONE;
        }
        if (active && node == B200_DONE) {
          if (best.entry < 0)
            finish_miss(idx);
          else
            store_hit(hits, idx, best.t, best.entry, S.entry_queue[best.entry]);
          active = false;
        }
      }
      continue;
    }
    while (active) {
      if (SPEC) {
        int32_t pending = B200_DONE;
        for (;;) {
          if (node < 0) { // a leaf or the end
            if (node == B200_DONE || pending != B200_DONE) break;
            pending = node;
            node = sp > 0 ? stack[--sp] : B200_DONE;
            continue;
          }
          node = WIDE ? node_step4(S, node, rs, lo32, hi32, stack, sp) : node_step(S, node, rs, lo32, hi32, stack, sp);
          if (node == B200_DONE && sp > 0) node = stack[--sp];
        }
        while (pending != B200_DONE) { // the postponed leaf, then whatever leaves follow directly
          leaf_test(S, pending, r, 0.001, B200_INF, st, bounce, best);
          if (best.entry >= 0) hi32 = __double2float_ru(best.t * rs.tsc);
          if (node < 0 && node != B200_DONE) {
            pending = node;
            node = sp > 0 ? stack[--sp] : B200_DONE;
          } else {
            pending = B200_DONE;
          }
        }
      } else {
        while (node >= 0) {
          node = WIDE ? node_step4(S, node, rs, lo32, hi32, stack, sp) : node_step(S, node, rs, lo32, hi32, stack, sp);
          if (node == B200_DONE && sp > 0) node = stack[--sp];
        }
        if (node != B200_DONE) {
          leaf_test(S, node, r, 0.001, B200_INF, st, bounce, best);
          if (best.entry >= 0) hi32 = __double2float_ru(best.t * rs.tsc);
          node = sp > 0 ? stack[--sp] : B200_DONE;
        }
      }
      if (node == B200_DONE) {
        if (best.entry < 0)
          finish_miss(idx);
        else
          store_hit(hits, idx, best.t, best.entry, S.entry_queue[best.entry]);
        active = false;
      } else if (count > 0 && __popc(__activemask()) < keep) {
        break;
      }
    }
  }
}

// Ray-reordering experiment (B200RT_OPT_RAY_SORT): one key per live path from its ray, so that paths with similar rays end up
// in the same 32-ray batch of the extend kernel.  mode bit 0: direction octant (3 bits); bit 1: 4 x 4 cells per face of the
// direction cube (2 + 4 bits: major axis, then the two minor components); bit 2: origin cell on a 16^3 grid over the scene
// bounds (12 bits, Morton order).  The most significant bits are the direction (rays of one direction class together, ordered
// by where they start).  Slots beyond n_cur get the largest key.
DEV uint32_t morton3_4(uint32_t x, uint32_t y, uint32_t z) { // 4 bits per axis
  uint32_t m = 0;
#pragma unroll
  for (int b = 0; b < 4; b++) m |= ((x >> b & 1u) << (3 * b)) | ((y >> b & 1u) << (3 * b + 1)) | ((z >> b & 1u) << (3 * b + 2));
  return m;
}
__global__ void __launch_bounds__(256) k_sort_keys(const WaveCtl *c, PathBuf cur, WaveParams wp, int mode, uint32_t *keys, uint32_t *vals, uint32_t capacity) {
  const uint32_t n = c->n_cur;
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < capacity; i += gridDim.x * blockDim.x) {
    uint32_t key = 0xffffffffu;
    if (i < n) {
      double2 q0, q1, q2, q3;
      ld256(quad_ptr(cur, 0, i), q0, q1);
      ld256(quad_ptr(cur, 2, i), q2, q3);
      const double ox = q0.x, oy = q0.y, oz = q1.x, dx = q1.y, dy = q2.x, dz = q2.y;
      key = 0;
      if (mode & 4) {
        auto cell = [](double o, double bound) { const double u = (o / bound + 1.0) * 8.0; return (uint32_t)(u < 0 ? 0 : (u > 15 ? 15 : (int)u)); };
        key = morton3_4(cell(ox, wp.bound[0]), cell(oy, wp.bound[1]), cell(oz, wp.bound[2]));
      }
      if (mode & 2) {
        const double ax = fabs(dx), ay = fabs(dy), az = fabs(dz);
        const int major = ax >= ay && ax >= az ? 0 : (ay >= az ? 1 : 2);
        const double m = major == 0 ? ax : (major == 1 ? ay : az), u = (major == 0 ? dy : dx) / m, v = (major == 2 ? dy : dz) / m;
        auto q4 = [](double t) { const int k = (int)((t + 1.0) * 2.0); return (uint32_t)(k < 0 ? 0 : (k > 3 ? 3 : k)); };
        key |= ((uint32_t)major << 4 | q4(u) << 2 | q4(v)) << 12;
      }
      if (mode & 1) key |= ((dx < 0 ? 1u : 0u) | (dy < 0 ? 2u : 0u) | (dz < 0 ? 4u : 0u)) << 18;
    }
    keys[i] = key;
    vals[i] = i;
  }
}

// Sorts path indices into the queue of what they hit.  Separate from the traversal so that it is perfectly
// coalesced.  Same-address global atomics run at about one per 2 ns on B200, so a warp-aggregated append (131 k
// atomics per 4 Mi-path wave on 3-4 hot counters) costs more than the traversal of a short wave; here a CTA ranks a
// chunk of 2048 paths in shared memory (ballot/popc per warp, one shared atomic per warp and queue) and claims its
// range of every queue with ONE global atomic per queue per chunk.
#define B200_CLASSIFY_K 8
__global__ void __launch_bounds__(256) k_classify(WaveCtl *c, const HitRef *hits, Queues Q) {
  __shared__ uint32_t s_count[B200_NQUEUES + 1], s_base[B200_NQUEUES]; // [B200_Q_MISS_DONE]: counted only
  const uint32_t n = c->n_cur;
  const int lane = threadIdx.x & 31;
  const uint32_t chunk = 256 * B200_CLASSIFY_K;
  for (uint32_t cbase = blockIdx.x * chunk; cbase < n; cbase += gridDim.x * chunk) {
    if (threadIdx.x <= B200_NQUEUES) s_count[threadIdx.x] = 0;
    __syncthreads();
    int which[B200_CLASSIFY_K];
    uint32_t rank[B200_CLASSIFY_K];
#pragma unroll
    for (int k = 0; k < B200_CLASSIFY_K; k++) {
      const uint32_t i = cbase + k * 256 + threadIdx.x;
      which[k] = i < n ? __ldg(reinterpret_cast<const int4 *>(hits + i)).w : -1;
      rank[k] = 0;
      {
        const unsigned m = __ballot_sync(0xffffffffu, which[k] == B200_Q_MISS_DONE);
        if (m != 0 && lane == 0) atomicAdd(&s_count[B200_Q_MISS_DONE], (uint32_t)__popc(m));
        if (which[k] == B200_Q_MISS_DONE) which[k] = -1;
      }
#pragma unroll
      for (int q = 0; q < B200_NQUEUES; q++) {
        const unsigned m = __ballot_sync(0xffffffffu, which[k] == q);
        if (m == 0) continue; // warp-uniform
        uint32_t b = 0;
        if (lane == __ffs(m) - 1) b = atomicAdd(&s_count[q], (uint32_t)__popc(m));
        b = __shfl_sync(0xffffffffu, b, __ffs(m) - 1);
        if (which[k] == q) rank[k] = b + (uint32_t)__popc(m & ((1u << lane) - 1u));
      }
    }
    __syncthreads();
    if (threadIdx.x < B200_NQUEUES) s_base[threadIdx.x] = s_count[threadIdx.x] ? atomicAdd(&c->q_count[threadIdx.x], s_count[threadIdx.x]) : 0;
    if (threadIdx.x == B200_NQUEUES && s_count[B200_NQUEUES]) atomicAdd(&c->q_count[B200_NQUEUES], s_count[B200_NQUEUES]);
    __syncthreads();
#pragma unroll
    for (int k = 0; k < B200_CLASSIFY_K; k++)
      if (which[k] >= 0) Q.q[which[k]][s_base[which[k]] + rank[k]] = cbase + k * 256 + threadIdx.x;
    __syncthreads();
  }
}

} // namespace b200
