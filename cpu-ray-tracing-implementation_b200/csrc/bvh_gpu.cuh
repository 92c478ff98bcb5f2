// bvh_gpu.cuh — the compacted BVH built ON the device: a linear BVH in the sense of Karras 2012.
//
// Stands in for bvh_node's constructor (src/bvh_node.h:18-47) like the host SAH builder (bvh_build.h) does, for scenes
// where build time matters (Sponza-sized meshes, animation): Morton codes of the box centroids -> radix sort -> one
// thread per inner node finds its range and split from the sorted codes -> one thread per leaf walks up and the second
// arrival at a node writes it (both children's boxes, rounded outward to binary32, in the 64-byte layout of
// scene_dev.cuh).  One primitive per leaf.  The tree is of lower quality than the SAH tree; which tree is traversed never
// changes a hit (ties go by Entry::order, traverse.cuh), so the parity tests run unchanged on either.
//
// The sort is cub::DeviceRadixSort (library plumbing, not on the per-ray path); every other kernel is written here.
#pragma once
#include <cub/device/device_radix_sort.cuh>

#include "flatten.h"
#include "scene_dev.cuh"

namespace b200 {

struct LbvhStats { // written by the kernels, read back once
  unsigned long long cbounds[6]; // centroid bounds as order-preserving keys: lo x,y,z then hi x,y,z
  unsigned int plane_max[3];     // max |finite fp32 plane| per axis (bit pattern of a non-negative float)
  unsigned int max_depth;
  double sah_acc;  // sum of 1.2 * area over inner nodes + area over leaves
  double root_area;
};

// order-preserving map double -> uint64 (for atomicMin / atomicMax)
__device__ inline unsigned long long lbvh_key(double x) {
  const unsigned long long b = (unsigned long long)__double_as_longlong(x);
  return (b >> 63) ? ~b : (b | 0x8000000000000000ull);
}
__device__ inline double lbvh_unkey(unsigned long long k) {
  const unsigned long long b = (k >> 63) ? (k & 0x7fffffffffffffffull) : ~k;
  return __longlong_as_double((long long)b);
}
__device__ inline double lbvh_centroid(const Box &b, int k) {
  const double c = 0.5 * (b.lo[k] + b.hi[k]);
  return isfinite(c) ? c : 0.0; // same rule as the host builder
}

__global__ void k_lbvh_init(LbvhStats *st) {
  for (int k = 0; k < 3; k++) st->cbounds[k] = ~0ull, st->cbounds[3 + k] = 0ull, st->plane_max[k] = 0u;
  st->max_depth = 0;
  st->sah_acc = 0;
  st->root_area = 0;
}

__global__ void __launch_bounds__(256) k_lbvh_bounds(const Box *boxes, uint32_t n, LbvhStats *st) {
  __shared__ unsigned long long s_lo[3][8], s_hi[3][8];
  unsigned long long lo[3] = {~0ull, ~0ull, ~0ull}, hi[3] = {0ull, 0ull, 0ull};
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const Box b = boxes[i];
    for (int k = 0; k < 3; k++) {
      const unsigned long long key = lbvh_key(lbvh_centroid(b, k));
      lo[k] = key < lo[k] ? key : lo[k];
      hi[k] = key > hi[k] ? key : hi[k];
    }
  }
  for (int k = 0; k < 3; k++) {
    for (int off = 16; off > 0; off >>= 1) {
      const unsigned long long a = __shfl_xor_sync(0xffffffffu, lo[k], off), b = __shfl_xor_sync(0xffffffffu, hi[k], off);
      lo[k] = a < lo[k] ? a : lo[k];
      hi[k] = b > hi[k] ? b : hi[k];
    }
    if ((threadIdx.x & 31) == 0) s_lo[k][threadIdx.x >> 5] = lo[k], s_hi[k][threadIdx.x >> 5] = hi[k];
  }
  __syncthreads();
  if (threadIdx.x < 3) {
    const int k = threadIdx.x;
    unsigned long long a = s_lo[k][0], b = s_hi[k][0];
    for (int w = 1; w < 8; w++) a = s_lo[k][w] < a ? s_lo[k][w] : a, b = s_hi[k][w] > b ? s_hi[k][w] : b;
    atomicMin(&st->cbounds[k], a);
    atomicMax(&st->cbounds[3 + k], b);
  }
}

__device__ inline uint32_t lbvh_spread10(uint32_t v) { // 10 bits -> every third bit
  v = (v * 0x00010001u) & 0xFF0000FFu;
  v = (v * 0x00000101u) & 0x0F00F00Fu;
  v = (v * 0x00000011u) & 0xC30C30C3u;
  v = (v * 0x00000005u) & 0x49249249u;
  return v;
}

__global__ void __launch_bounds__(256) k_lbvh_morton(const Box *boxes, uint32_t n, const LbvhStats *st, uint32_t *codes, uint32_t *ids) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const Box b = boxes[i];
  uint32_t q[3];
  for (int k = 0; k < 3; k++) {
    const double lo = lbvh_unkey(st->cbounds[k]), hi = lbvh_unkey(st->cbounds[3 + k]);
    const double ext = hi - lo;
    double u = ext > 0 ? (lbvh_centroid(b, k) - lo) / ext : 0.0;
    u = u < 0 ? 0 : (u > 1 ? 1 : u);
    const int v = (int)(u * 1024.0);
    q[k] = (uint32_t)(v > 1023 ? 1023 : v);
  }
  codes[i] = (lbvh_spread10(q[0]) << 2) | (lbvh_spread10(q[1]) << 1) | lbvh_spread10(q[2]);
  ids[i] = i;
}

// length of the common prefix of the keys (code, position) of sorted positions i and j; -1 outside the array
__device__ inline int lbvh_delta(const uint32_t *codes, int n, int i, int j) {
  if (j < 0 || j >= n) return -1;
  const uint32_t a = codes[i], b = codes[j];
  return a != b ? __clz(a ^ b) : 32 + __clz((uint32_t)i ^ (uint32_t)j);
}

// Child references while building: >= 0 inner node, < 0 leaf at sorted position ~ref.
__global__ void __launch_bounds__(256) k_lbvh_hierarchy(const uint32_t *codes, int n, int2 *children, int *parent_inner, int *parent_leaf) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n - 1) return;
  const int d = lbvh_delta(codes, n, i, i + 1) - lbvh_delta(codes, n, i, i - 1) >= 0 ? 1 : -1;
  const int dmin = lbvh_delta(codes, n, i, i - d);
  int lmax = 2;
  while (lbvh_delta(codes, n, i, i + lmax * d) > dmin) lmax <<= 1;
  int l = 0;
  for (int t = lmax >> 1; t >= 1; t >>= 1)
    if (lbvh_delta(codes, n, i, i + (l + t) * d) > dmin) l += t;
  const int j = i + l * d;
  const int dnode = lbvh_delta(codes, n, i, j);
  int s = 0;
  for (int t = (l + 1) >> 1;; t = (t + 1) >> 1) { // ceil(l / 2), ceil(l / 4), ..., 1
    if (lbvh_delta(codes, n, i, i + (s + t) * d) > dnode) s += t;
    if (t == 1) break;
  }
  const int gamma = i + s * d + (d < 0 ? -1 : 0);
  const int lo = i < j ? i : j, hi = i < j ? j : i;
  const int left = lo == gamma ? ~gamma : gamma;
  const int right = hi == gamma + 1 ? ~(gamma + 1) : gamma + 1;
  children[i] = make_int2(left, right);
  if (left >= 0) parent_inner[left] = i; else parent_leaf[~left] = i;
  if (right >= 0) parent_inner[right] = i; else parent_leaf[~right] = i;
  if (i == 0) parent_inner[0] = -1;
}

struct Box32 {
  float v[6]; // x0 x1 y0 y1 z0 z1, rounded outward
};
__device__ inline Box32 lbvh_round(const Box &b) {
  Box32 r;
  for (int k = 0; k < 3; k++) r.v[2 * k] = __double2float_rd(b.lo[k]), r.v[2 * k + 1] = __double2float_ru(b.hi[k]);
  return r;
}
__device__ inline Box32 lbvh_union(const Box32 &a, const Box32 &b) {
  Box32 r;
  for (int k = 0; k < 3; k++) r.v[2 * k] = fminf(a.v[2 * k], b.v[2 * k]), r.v[2 * k + 1] = fmaxf(a.v[2 * k + 1], b.v[2 * k + 1]);
  return r;
}
__device__ inline double lbvh_area(const Box32 &b) {
  const double dx = (double)b.v[1] - b.v[0], dy = (double)b.v[3] - b.v[2], dz = (double)b.v[5] - b.v[4];
  if (!(dx >= 0 && dy >= 0 && dz >= 0)) return 0;
  const double a = 2 * (dx * dy + dy * dz + dz * dx);
  return isfinite(a) ? a : 1e300;
}

// One thread per leaf walks towards the root; the second thread to arrive at an inner node owns it: it reads both
// children's boxes (its own in registers, the sibling's from `node_box` or straight from the primitive), writes the
// finished 64-byte node and carries the union upwards.
__global__ void __launch_bounds__(256) k_lbvh_refit(const Box *boxes, const uint32_t *ids, int n, const int2 *children, const int *parent_inner,
                                                     const int *parent_leaf, unsigned int *arrived, Box32 *node_box, BvhNode *nodes, LbvhStats *st) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= n) return;
  Box32 mine = lbvh_round(boxes[ids[j]]);
  for (int k = 0; k < 6; k++) {
    const float a = fabsf(mine.v[k]);
    if (isfinite(a)) atomicMax(&st->plane_max[k >> 1], __float_as_uint(a));
  }
  double sah = lbvh_area(mine);
  int me = ~j, p = parent_leaf[j];
  if (p < 0) st->root_area = sah; // a one-primitive world: the leaf is the root
  while (p >= 0) {
    __threadfence();
    if (atomicAdd(&arrived[p], 1u) == 0u) break; // the sibling subtree is not finished: its thread will come by later
    __threadfence();
    const int2 ch = children[p];
    const int sib = ch.x == me ? ch.y : ch.x;
    Box32 other;
    if (sib < 0) {
      other = lbvh_round(boxes[ids[~sib]]);
    } else {
      const volatile float *q = node_box[sib].v;
      for (int k = 0; k < 6; k++) other.v[k] = q[k];
    }
    const Box32 &lb = ch.x == me ? mine : other, &rb = ch.x == me ? other : mine;
    BvhNode nd;
    for (int k = 0; k < 6; k++) nd.lbox[k] = lb.v[k], nd.rbox[k] = rb.v[k];
    nd.left = ch.x >= 0 ? ch.x : ~(int32_t)((uint32_t)(~ch.x) << 3); // leaf: first entry = sorted position, count - 1 = 0
    nd.right = ch.y >= 0 ? ch.y : ~(int32_t)((uint32_t)(~ch.y) << 3);
    nd.pad[0] = nd.pad[1] = 0;
    nodes[p] = nd;
    mine = lbvh_union(mine, other);
    for (int k = 0; k < 6; k++) node_box[p].v[k] = mine.v[k];
    sah += 1.2 * lbvh_area(mine);
    if (p == 0) st->root_area = lbvh_area(mine);
    me = p;
    p = parent_inner[p];
  }
  atomicAdd(&st->sah_acc, sah);
  // depth of this leaf = inner nodes above it + 1, whoever wrote them
  unsigned int dd = 1;
  for (int q = parent_leaf[j]; q >= 0; q = parent_inner[q]) dd++;
  atomicMax(&st->max_depth, dd);
}

__global__ void __launch_bounds__(256) k_lbvh_gather(const Entry *in, const uint32_t *ids, uint32_t n, Entry *out) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = in[ids[i]];
}

} // namespace b200
