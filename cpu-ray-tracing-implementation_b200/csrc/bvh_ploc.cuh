// bvh_ploc.cuh — the BVH2 topology built on the device by parallel locally-ordered clustering (PLOC, Meister & Bittner 2018),
// the quality half of SURVEY.md §8f rank 1: an agglomerative build whose trees come close to the host's binned-SAH trees
// (bvh_build.h) at a fraction of the build time.  Like every builder here it stands in for bvh_node's constructor
// (src/bvh_node.h:18-47) and cannot change a hit: ties go by Entry::order (traverse.cuh).
//
//   clusters = the primitives in Morton order (the codes and the radix sort of bvh_gpu.cuh)
//   repeat until one cluster is left:
//     k_ploc_nearest : every cluster looks R positions to either side for the partner whose union box has the smallest area
//     k_ploc_flag    : a pair that chose each other merges; the lower index keeps the slot, the higher one disappears
//     (exclusive scan of {survivors, merges}: cub, library plumbing)
//     k_ploc_merge   : survivors move to their scanned slot; a merging pair writes one 64-byte node (both children's boxes, rounded
//                      outward to binary32 once, at the leaves; unions of binary32 boxes are exact) and carries the union on
//
// Node indices come from the scan, not from atomics: the same scene gives the same tree, children always precede their parent and
// the root is the last node.  Leaves hold one primitive here; the host pass bvh_build.h:finish_tree() then merges leaf pairs by
// the same SAH rule as the host builder, lays the nodes out in depth-first order and hands the tree to the BVH4 collapse.
#pragma once
#include <cub/device/device_scan.cuh>

#include "bvh_gpu.cuh"

namespace b200 {

#define B200_PLOC_RADIUS 4 // positions searched to either side; measured 2 ... 128 (profiles/r02_ap, r02_aq): small radii give the faster trees here

struct __align__(16) PlocCluster {
  float v[6];  // x0 x1 y0 y1 z0 z1
  int32_t ref; // >= 0 inner node, < 0 leaf reference (traverse.cuh encoding, one primitive)
  uint32_t pad;
};

__device__ inline PlocCluster ploc_load(const PlocCluster *p) {
  const float4 a = __ldg(reinterpret_cast<const float4 *>(p)), b = __ldg(reinterpret_cast<const float4 *>(p) + 1);
  PlocCluster c;
  c.v[0] = a.x, c.v[1] = a.y, c.v[2] = a.z, c.v[3] = a.w, c.v[4] = b.x, c.v[5] = b.y;
  c.ref = __float_as_int(b.z), c.pad = 0;
  return c;
}
// area of the union box; symmetric in its arguments bit for bit (fminf / fmaxf are), which mutual choices rely on
__device__ inline double ploc_union_area(const PlocCluster &a, const PlocCluster &b) {
  Box32 u;
  for (int k = 0; k < 3; k++) u.v[2 * k] = fminf(a.v[2 * k], b.v[2 * k]), u.v[2 * k + 1] = fmaxf(a.v[2 * k + 1], b.v[2 * k + 1]);
  return lbvh_area(u);
}

__global__ void __launch_bounds__(256) k_ploc_init(const Box *boxes, const uint32_t *ids, uint32_t n, PlocCluster *out, LbvhStats *st) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const Box32 b = lbvh_round(boxes[ids[i]]);
  PlocCluster c;
  for (int k = 0; k < 6; k++) {
    c.v[k] = b.v[k];
    const float a = fabsf(b.v[k]);
    if (isfinite(a)) atomicMax(&st->plane_max[k >> 1], __float_as_uint(a));
  }
  c.ref = ~(int32_t)(i << 3); // leaf: first entry = sorted position, count - 1 = 0
  c.pad = 0;
  out[i] = c;
}

// nn[i] = the position within R of i whose union with i has the smallest area; ties go to the smaller position, which guarantees
// that the pair achieving the global minimum (lowest positions among equals) chooses each other: every round merges something.
__global__ void __launch_bounds__(256) k_ploc_nearest(const PlocCluster *cl, int m, int radius, int *nn) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= m) return;
  const PlocCluster a = ploc_load(cl + i);
  const int lo = i - radius < 0 ? 0 : i - radius, hi = i + radius > m - 1 ? m - 1 : i + radius;
  double best = 1.7976931348623157e308;
  int bj = -1;
  for (int j = lo; j <= hi; j++) {
    if (j == i) continue;
    const double ar = ploc_union_area(a, ploc_load(cl + j));
    if (bj < 0 || ar < best) best = ar, bj = j;
  }
  nn[i] = bj;
}

// low word: 1 if position i survives the round (alone or as the lower half of a merging pair); high word: 1 if it merges
__global__ void __launch_bounds__(256) k_ploc_flag(const int *nn, int m, unsigned long long *flags) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= m) return;
  const int j = nn[i];
  const bool mutual = j >= 0 && nn[j] == i;
  flags[i] = (mutual && i > j) ? 0ull : (1ull | (mutual ? 1ull << 32 : 0ull));
}

__global__ void __launch_bounds__(256) k_ploc_merge(const PlocCluster *cl, const int *nn, const unsigned long long *flags, const unsigned long long *scan, int m,
                                                    uint32_t node_base, PlocCluster *out, BvhNode *nodes, unsigned int *totals) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= m) return;
  const unsigned long long f = flags[i], s = scan[i];
  if (i == m - 1) totals[0] = (unsigned int)(s + f), totals[1] = (unsigned int)((s + f) >> 32); // survivors, merges of this round
  if (!(f & 1ull)) return;
  const uint32_t dst = (uint32_t)s;
  PlocCluster a = ploc_load(cl + i);
  if (f >> 32) {
    const PlocCluster b = ploc_load(cl + nn[i]);
    const uint32_t node = node_base + (uint32_t)(s >> 32);
    BvhNode nd;
    for (int k = 0; k < 6; k++) nd.lbox[k] = a.v[k], nd.rbox[k] = b.v[k];
    nd.left = a.ref, nd.right = b.ref;
    nd.pad[0] = nd.pad[1] = 0;
    nodes[node] = nd;
    for (int k = 0; k < 3; k++) a.v[2 * k] = fminf(a.v[2 * k], b.v[2 * k]), a.v[2 * k + 1] = fmaxf(a.v[2 * k + 1], b.v[2 * k + 1]);
    a.ref = (int32_t)node;
  }
  out[dst] = a;
}

} // namespace b200
