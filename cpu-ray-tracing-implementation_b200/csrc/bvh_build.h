// bvh_build.h — host builder of the compacted BVH2 that stands in for the reference's bvh_node tree.
//
// The reference splits at the median of the objects sorted by bounding-box x-min, always on axis 0
// (src/bvh_node.h:20-31, 42-44), which costs ~75 node visits per ray on a field of spheres.
// Here: binned surface-area heuristic over all three axes, leaves of up to `max_leaf` entries, nodes that
// carry both children's boxes so one 128-byte fetch decides both sides.  Which tree is traversed does not
// change the answer: ties are resolved by Entry::order (flatten.h).
#pragma once
#include <chrono>
#include <cmath>
#include <numeric>
#include <stdexcept>
#include <thread>

#include "flatten.h"

namespace b200 {

struct BuiltBvh {
  std::vector<BvhNode> nodes;
  std::vector<uint32_t> perm; // perm[k] = index into the flattener's entry order of the k-th entry in leaf order
  int32_t root = B200_EMPTY_ROOT;
  uint32_t n_leaves = 0, max_depth = 0;
  double sah_cost = 0, build_ms = 0;
  double bound[3] = {0, 0, 0}; // max |finite fp32 plane| per axis
};

class BvhBuilder {
public:
  // threads: 0 = as many as the machine has (large scenes only: subtrees of >= kParallelMin primitives in the top kParallelDepth
  // levels are built by their own threads), 1 = serial.  The tree does not depend on it: node indices, boxes and the entry
  // permutation are identical (tests/test_emu_device_functions.py), only the SAH cost's last bits (summation order) may differ.
  BvhBuilder(const std::vector<Box> &boxes, int max_leaf, bool peel = true, bool sweep = false, bool area_axis = false, int threads = 0)
      : boxes_(boxes), max_leaf_(std::max(1, std::min(8, max_leaf))), peel_(peel), sweep_(sweep), area_axis_(area_axis),
        parallel_(threads == 0 ? std::thread::hardware_concurrency() > 1 : threads > 1) {}

  BuiltBvh run() {
    auto t0 = std::chrono::steady_clock::now();
    const size_t n = boxes_.size();
    out_.perm.resize(n);
    std::iota(out_.perm.begin(), out_.perm.end(), 0u);
    cent_.resize(n * 3);
    area_.resize(n);
    larea_.resize(n);
    for (size_t i = 0; i < n; i++) {
      for (int k = 0; k < 3; k++) {
        double c = 0.5 * (boxes_[i].lo[k] + boxes_[i].hi[k]);
        cent_[3 * i + k] = std::isfinite(c) ? c : 0.0;
      }
      const double a = boxes_[i].area();
      area_[i] = std::isfinite(a) ? a : 1e300;
      larea_[i] = area_[i] > 1e-300 ? std::log2(area_[i]) : -1000.0;
    }
    if (n > 0) {
      Box rb;
      Acc acc;
      out_.root = build(acc, 0, (uint32_t)n, 1, rb);
      out_.nodes = std::move(acc.nodes);
      out_.n_leaves = acc.n_leaves, out_.max_depth = acc.max_depth;
      for (int k = 0; k < 3; k++) out_.bound[k] = acc.bound[k];
      root_area_ = rb.area();
      if (root_area_ > 0) out_.sah_cost = acc.cost / root_area_;
    }
    out_.build_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    return std::move(out_);
  }

private:
  static constexpr int kBins = 16;
  static constexpr uint32_t kSweep = 32;
  const std::vector<Box> &boxes_;
  int max_leaf_;
  bool peel_, sweep_, area_axis_;
  std::vector<double> cent_, area_, larea_;
  double key(uint32_t id, int axis) const { return axis < 3 ? cent_[3 * id + axis] : larea_[id]; } // axis 3: log2 of the box area
  bool parallel_;
  static constexpr uint32_t kParallelMin = 16384, kParallelDepth = 5;
  BuiltBvh out_;
  double root_area_ = 0;
  struct Acc { // what one (sub)tree build produces besides its share of out_.perm
    std::vector<BvhNode> nodes; // indices local to this vector
    double cost = 0;
    uint32_t n_leaves = 0, max_depth = 0;
    double bound[3] = {0, 0, 0};
  };

  static float round_down(double x) {
    float f = (float)x;
    if ((double)f > x) f = std::nextafterf(f, -INFINITY);
    return f;
  }
  static float round_up(double x) {
    float f = (float)x;
    if ((double)f < x) f = std::nextafterf(f, INFINITY);
    return f;
  }
  static void set_box(Acc &acc, float dst[6], const Box &b) {
    for (int k = 0; k < 3; k++) {
      dst[2 * k] = round_down(b.lo[k]);
      dst[2 * k + 1] = round_up(b.hi[k]);
      for (int e = 0; e < 2; e++) {
        const double a = std::fabs((double)dst[2 * k + e]);
        if (std::isfinite(a) && a > acc.bound[k]) acc.bound[k] = a;
      }
    }
  }

  static int32_t make_leaf(Acc &acc, uint32_t begin, uint32_t end, uint32_t depth, const Box &b) {
    acc.n_leaves++;
    acc.max_depth = std::max(acc.max_depth, depth);
    acc.cost += b.area() * (end - begin);
    return ~(int32_t)((begin << 3) | (end - begin - 1));
  }

  int32_t build(Acc &acc, uint32_t begin, uint32_t end, uint32_t depth, Box &bounds) {
    const uint32_t count = end - begin;
    bounds.reset();
    Box cb;
    cb.reset();
    uint32_t big = begin; // the primitive with the largest box (the peel-off candidate below)
    double big_area = -1, alo = INFINITY, ahi = -INFINITY;
    for (uint32_t i = begin; i < end; i++) {
      const uint32_t id = out_.perm[i];
      bounds.add(boxes_[id]);
      cb.add(&cent_[3 * id]);
      if (area_[id] > big_area) big_area = area_[id], big = i;
      alo = std::min(alo, larea_[id]), ahi = std::max(ahi, larea_[id]);
    }
    if (count == 1) return make_leaf(acc, begin, end, depth, bounds);

    // best binned SAH split
    double best_cost = INFINITY;
    int best_axis = -1, best_plane = -1;
    if (depth < 40) {
      // axes 0..2: centroid coordinates.  Axis 3 (optional): log2 of the box area, i.e. "the large primitives against the small ones", a
      // partition no centroid plane expresses (walls around furniture); only where the areas in the node span more than a factor 4
      // Measured (profiles/r02_aw_sah_area_axis.jsonl): Sponza stand-in +1.1 %, Cornell +0.9 %, RTNW -0.9 %, build time +8 %: off.
      for (int axis = 0; axis < (area_axis_ ? 4 : 3); axis++) {
        const double lo = axis < 3 ? cb.lo[axis] : alo, ext = axis < 3 ? cb.hi[axis] - cb.lo[axis] : ahi - alo;
        if (!(ext > (axis < 3 ? 0.0 : 2.0)) || !std::isfinite(ext)) continue;
        Box bin_box[kBins];
        uint32_t bin_cnt[kBins] = {0};
        for (auto &b : bin_box) b.reset();
        const double scale = kBins / ext;
        for (uint32_t i = begin; i < end; i++) {
          const uint32_t id = out_.perm[i];
          int b = (int)((key(id, axis) - lo) * scale);
          b = b < 0 ? 0 : (b >= kBins ? kBins - 1 : b);
          bin_cnt[b]++;
          bin_box[b].add(boxes_[id]);
        }
        double right_area[kBins];
        uint32_t right_cnt[kBins];
        Box acc;
        acc.reset();
        uint32_t c = 0;
        for (int b = kBins - 1; b > 0; b--) {
          if (bin_cnt[b]) acc.add(bin_box[b]);
          c += bin_cnt[b];
          right_area[b] = acc.area();
          right_cnt[b] = c;
        }
        acc.reset();
        c = 0;
        for (int b = 0; b < kBins - 1; b++) {
          if (bin_cnt[b]) acc.add(bin_box[b]);
          c += bin_cnt[b];
          if (c == 0 || right_cnt[b + 1] == 0) continue;
          double la = acc.area(), ra = right_area[b + 1];
          if (!std::isfinite(la)) la = 1e300;
          if (!std::isfinite(ra)) ra = 1e300;
          const double cost = la * c + ra * right_cnt[b + 1];
          if (cost < best_cost) best_cost = cost, best_axis = axis, best_plane = b;
        }
      }
    }
    // Small nodes, optional: the exact sweep over every split of the centroid order on every axis instead of the 16 bins.  Measured
    // (profiles/r02_ar_sah_sweep.jsonl): +50 % build time, traversal within noise on all five configs, so it is off.
    int sweep_axis = -1;
    uint32_t sweep_left = 0;
    if (sweep_ && depth < 40 && count <= kSweep) {
      uint32_t order[kSweep];
      double right_area[kSweep];
      for (int axis = 0; axis < 3; axis++) {
        for (uint32_t i = 0; i < count; i++) order[i] = out_.perm[begin + i];
        std::stable_sort(order, order + count, [&](uint32_t a, uint32_t b) { return cent_[3 * a + axis] < cent_[3 * b + axis]; });
        Box acc;
        acc.reset();
        for (uint32_t i = count - 1; i > 0; i--) {
          acc.add(boxes_[order[i]]);
          const double a = acc.area();
          right_area[i] = std::isfinite(a) ? a : 1e300;
        }
        acc.reset();
        for (uint32_t i = 0; i + 1 < count; i++) {
          acc.add(boxes_[order[i]]);
          double la = acc.area();
          if (!std::isfinite(la)) la = 1e300;
          const double cost = la * (i + 1) + right_area[i + 1] * (count - i - 1);
          if (cost < best_cost) best_cost = cost, sweep_axis = axis, sweep_left = i + 1;
        }
      }
    }
    // One more candidate that centroid bins cannot express: the primitive with the largest box on its own against all the others.
    // A primitive that spans the node (a fog boundary, a ground sphere, a wall) otherwise follows the centroid splits downwards and
    // makes every node on its way as large as itself.  Measured (profiles/r02_aq_sah_peel_ploc_radius.jsonl): RTNW's tree cost 7.43 ->
    // 2.66 and its frame +17.7 % (3 188 -> 3 754 Mrays/s); the other configs unchanged.
    int32_t peel = -1;
    if (peel_ && depth < 40 && count > 2 && big_area >= 0.25 * bounds.area()) { // (a mesh triangle never gets here: no second pass for it)
      Box rest;
      rest.reset();
      for (uint32_t i = begin; i < end; i++)
        if (i != big) rest.add(boxes_[out_.perm[i]]);
      double ra = rest.area();
      if (!std::isfinite(ra)) ra = 1e300;
      const double cost = big_area + ra * (count - 1);
      if (cost < best_cost) best_cost = cost, peel = (int32_t)big;
    }
    double parent_area = bounds.area();
    if (!std::isfinite(parent_area)) parent_area = 1e300;
    if (count <= (uint32_t)max_leaf_) {
      // leaf cost: count intersections; split cost: one node (two boxes, ~1.2 intersections' worth) + children
      const double leaf_cost = (double)count * parent_area;
      const double split_cost = (best_axis < 0 && peel < 0 && sweep_axis < 0) ? INFINITY : 1.2 * parent_area + best_cost;
      if (leaf_cost <= split_cost) return make_leaf(acc, begin, end, depth, bounds);
    }
    uint32_t mid;
    if (peel >= 0) {
      std::swap(out_.perm[begin], out_.perm[peel]);
      mid = begin + 1;
    } else if (sweep_axis >= 0) {
      std::stable_sort(out_.perm.begin() + begin, out_.perm.begin() + end, [&](uint32_t a, uint32_t b) { return cent_[3 * a + sweep_axis] < cent_[3 * b + sweep_axis]; });
      mid = begin + sweep_left;
    } else if (best_axis >= 0) {
      const double lo = best_axis < 3 ? cb.lo[best_axis] : alo, scale = kBins / (best_axis < 3 ? cb.hi[best_axis] - cb.lo[best_axis] : ahi - alo);
      auto it = std::partition(out_.perm.begin() + begin, out_.perm.begin() + end, [&](uint32_t id) {
        int b = (int)((key(id, best_axis) - lo) * scale);
        b = b < 0 ? 0 : (b >= kBins ? kBins - 1 : b);
        return b <= best_plane;
      });
      mid = (uint32_t)(it - out_.perm.begin());
    } else {
      mid = begin;
    }
    if (mid == begin || mid == end) { // degenerate (coincident centroids) or depth guard: median on the widest axis
      int axis = 0;
      for (int k = 1; k < 3; k++)
        if (cb.hi[k] - cb.lo[k] > cb.hi[axis] - cb.lo[axis]) axis = k;
      mid = begin + count / 2;
      std::nth_element(out_.perm.begin() + begin, out_.perm.begin() + mid, out_.perm.begin() + end,
                       [&](uint32_t a, uint32_t b) { return cent_[3 * a + axis] < cent_[3 * b + axis]; });
    }
    const uint32_t me = (uint32_t)acc.nodes.size();
    acc.nodes.emplace_back();
    std::memset(&acc.nodes[me], 0, sizeof(BvhNode));
    acc.cost += 1.2 * parent_area;
    Box lb, rb;
    int32_t l, r;
    if (parallel_ && depth <= kParallelDepth && mid - begin >= kParallelMin && end - mid >= kParallelMin) {
      // the two halves work on disjoint ranges of out_.perm and read everything else: the left one gets its own thread, both
      // build into their own node vectors, which are then appended in the order the serial recursion would have produced
      Acc la, ra;
      std::thread left([&] { l = build(la, begin, mid, depth + 1, lb); });
      r = build(ra, mid, end, depth + 1, rb);
      left.join();
      l = splice(acc, la, l);
      r = splice(acc, ra, r);
    } else {
      l = build(acc, begin, mid, depth + 1, lb);
      r = build(acc, mid, end, depth + 1, rb);
    }
    BvhNode &nd = acc.nodes[me];
    nd.left = l;
    nd.right = r;
    set_box(acc, nd.lbox, lb);
    set_box(acc, nd.rbox, rb);
    return (int32_t)me;
  }

  // appends a finished subtree to `acc`; returns its root reference in acc's numbering
  static int32_t splice(Acc &acc, Acc &sub, int32_t root) {
    const int32_t off = (int32_t)acc.nodes.size();
    for (BvhNode nd : sub.nodes) {
      if (nd.left >= 0) nd.left += off;
      if (nd.right >= 0) nd.right += off;
      acc.nodes.push_back(nd);
    }
    acc.cost += sub.cost;
    acc.n_leaves += sub.n_leaves;
    acc.max_depth = std::max(acc.max_depth, sub.max_depth);
    for (int k = 0; k < 3; k++) acc.bound[k] = std::max(acc.bound[k], sub.bound[k]);
    return root >= 0 ? root + off : root;
  }
};

// Host pass over a BVH2 TOPOLOGY that came from somewhere else (the device PLOC build, bvh_ploc.cuh): `in` holds one-primitive
// leaves (reference ~(position << 3)) and inner nodes whose children always precede them, the root last.  Produces what
// BvhBuilder::run() produces: leaves of up to max_leaf entries where the SAH rule of the host builder prefers a leaf to a split
// (count * A <= 1.2 * A + sum of child costs), nodes in depth-first order, perm[k] = input position of the k-th entry in leaf order.
inline BuiltBvh finish_tree(const std::vector<BvhNode> &in, uint32_t n_prims, int max_leaf, const double bound[3]) {
  auto t0 = std::chrono::steady_clock::now();
  BuiltBvh out;
  max_leaf = std::max(1, std::min(8, max_leaf));
  for (int k = 0; k < 3; k++) out.bound[k] = bound[k];
  out.perm.reserve(n_prims);
  if (n_prims == 0) return out;
  if (in.empty()) { // one primitive: the leaf is the root
    out.perm.push_back(0);
    out.root = ~(int32_t)0;
    out.n_leaves = 1, out.max_depth = 1, out.sah_cost = 1;
    return out;
  }
  auto area = [](const float *b) {
    const double dx = (double)b[1] - b[0], dy = (double)b[3] - b[2], dz = (double)b[5] - b[4];
    if (!(dx >= 0 && dy >= 0 && dz >= 0)) return 0.0;
    const double a = 2 * (dx * dy + dy * dz + dz * dx);
    return std::isfinite(a) ? a : 1e300;
  };
  auto union_area = [&](const BvhNode &nd) {
    float u[6];
    for (int k = 0; k < 3; k++) u[2 * k] = std::min(nd.lbox[2 * k], nd.rbox[2 * k]), u[2 * k + 1] = std::max(nd.lbox[2 * k + 1], nd.rbox[2 * k + 1]);
    return area(u);
  };
  // pass 1 (children precede parents): how many primitives a node holds if it becomes a leaf, 0 if it stays an inner node
  const size_t nn = in.size();
  std::vector<uint8_t> leaf_count(nn, 0);
  auto count_of = [&](int32_t ref) -> uint32_t { return ref < 0 ? (((uint32_t)~ref) & 7u) + 1u : leaf_count[ref]; };
  for (size_t i = 0; i < nn; i++) {
    const BvhNode &nd = in[i];
    if ((nd.left >= 0 && (size_t)nd.left >= i) || (nd.right >= 0 && (size_t)nd.right >= i)) throw std::runtime_error("finish_tree: a child does not precede its parent");
    const uint32_t cl = count_of(nd.left), cr = count_of(nd.right);
    if (cl == 0 || cr == 0 || cl + cr > (uint32_t)max_leaf) continue;
    const double A = union_area(nd), la = area(nd.lbox), ra = area(nd.rbox);
    if ((double)(cl + cr) * A <= 1.2 * A + la * cl + ra * cr) leaf_count[i] = (uint8_t)(cl + cr);
  }
  // pass 2: depth-first emission from the root (the last node)
  struct Job {
    int32_t node;  // input index
    int32_t slot;  // output index, already reserved
    uint32_t depth;
  };
  auto emit_leaf = [&](int32_t ref, uint32_t depth, double a) -> int32_t { // ref: an input leaf or an input node with leaf_count > 0
    const uint32_t begin = (uint32_t)out.perm.size();
    std::vector<int32_t> todo{ref};
    while (!todo.empty()) {
      const int32_t r = todo.back();
      todo.pop_back();
      if (r < 0)
        out.perm.push_back(((uint32_t)~r) >> 3);
      else
        todo.push_back(in[r].right), todo.push_back(in[r].left);
    }
    const uint32_t count = (uint32_t)out.perm.size() - begin;
    out.n_leaves++;
    out.max_depth = std::max(out.max_depth, depth);
    out.sah_cost += a * count;
    return ~(int32_t)((begin << 3) | (count - 1));
  };
  const int32_t root = (int32_t)nn - 1;
  if (leaf_count[root]) { // the whole scene fits one leaf
    out.root = emit_leaf(root, 1, union_area(in[root]));
    out.sah_cost = (double)leaf_count[root];
    return out;
  }
  std::vector<Job> stack;
  out.nodes.reserve(nn);
  out.nodes.emplace_back();
  stack.push_back({root, 0, 1});
  const double root_area = union_area(in[root]);
  while (!stack.empty()) {
    const Job j = stack.back();
    stack.pop_back();
    BvhNode nd = in[j.node];
    out.sah_cost += 1.2 * union_area(nd);
    const int32_t kids[2] = {nd.left, nd.right};
    int32_t refs[2];
    int32_t pending[2] = {-1, -1};
    // the left subtree follows its parent immediately (depth-first), so its slot is only known once it is emitted: reserve lazily
    for (int k = 0; k < 2; k++) {
      if (kids[k] < 0 || leaf_count[kids[k]]) {
        refs[k] = 0; // filled below, in child order, so that leaf entries appear left to right
        pending[k] = -2;
      }
    }
    // leaves first in child order, then inner children get their slots (right pushed first so that left is processed next)
    for (int k = 0; k < 2; k++)
      if (pending[k] == -2) refs[k] = emit_leaf(kids[k], j.depth + 1, area(k == 0 ? nd.lbox : nd.rbox));
    for (int k = 0; k < 2; k++)
      if (pending[k] != -2) {
        refs[k] = (int32_t)out.nodes.size();
        out.nodes.emplace_back();
      }
    if (pending[1] != -2) stack.push_back({kids[1], refs[1], j.depth + 1});
    if (pending[0] != -2) stack.push_back({kids[0], refs[0], j.depth + 1});
    nd.left = refs[0], nd.right = refs[1];
    nd.pad[0] = nd.pad[1] = 0;
    out.nodes[j.slot] = nd;
    out.max_depth = std::max(out.max_depth, j.depth);
  }
  out.root = 0;
  if (root_area > 0) out.sah_cost /= root_area;
  out.build_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
  return out;
}

// Small scenes only (O(n^3)): exact greedy agglomerative clustering (Walter et al. 2008): the pair of clusters whose union box has the
// smallest area merges, until one cluster is left; then finish_tree.  On a handful of large, overlapping primitives (a room with
// furniture) top-down centroid splits cannot separate "the walls" from "what stands inside"; bottom-up merging does.  The caller keeps
// whichever of this tree and the SAH builder's has the lower SAH cost.
inline BuiltBvh build_agglomerative(const std::vector<Box> &boxes, int max_leaf) {
  auto t0 = std::chrono::steady_clock::now();
  struct Cl {
    float v[6];
    int32_t ref;
  };
  const size_t n = boxes.size();
  std::vector<Cl> cl(n);
  double bound[3] = {0, 0, 0};
  for (size_t i = 0; i < n; i++) {
    for (int k = 0; k < 3; k++) {
      float lo = (float)boxes[i].lo[k], hi = (float)boxes[i].hi[k];
      if ((double)lo > boxes[i].lo[k]) lo = std::nextafterf(lo, -INFINITY);
      if ((double)hi < boxes[i].hi[k]) hi = std::nextafterf(hi, INFINITY);
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
  topo.reserve(n);
  while (cl.size() > 1) {
    size_t bi = 0, bj = 1;
    double best = INFINITY;
    for (size_t i = 0; i < cl.size(); i++)
      for (size_t j = i + 1; j < cl.size(); j++) {
        const double ar = union_area(cl[i], cl[j]);
        if (ar < best) best = ar, bi = i, bj = j;
      }
    BvhNode nd;
    std::memset(&nd, 0, sizeof nd);
    for (int k = 0; k < 6; k++) nd.lbox[k] = cl[bi].v[k], nd.rbox[k] = cl[bj].v[k];
    nd.left = cl[bi].ref, nd.right = cl[bj].ref;
    for (int k = 0; k < 3; k++) cl[bi].v[2 * k] = std::min(cl[bi].v[2 * k], cl[bj].v[2 * k]), cl[bi].v[2 * k + 1] = std::max(cl[bi].v[2 * k + 1], cl[bj].v[2 * k + 1]);
    cl[bi].ref = (int32_t)topo.size();
    topo.push_back(nd);
    cl.erase(cl.begin() + (long)bj);
  }
  BuiltBvh out = finish_tree(topo, (uint32_t)n, max_leaf, bound);
  out.build_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
  return out;
}

// BVH2 -> BVH4: every kept node adopts its grandchildren, largest box first, until it has four children or only leaves.
// Child references keep their encoding (inner: index into the NEW array; leaf: unchanged).
inline std::vector<BvhNode4> collapse4(const BuiltBvh &b2, int32_t &root4, uint32_t &max_depth4) {
  std::vector<BvhNode4> out;
  max_depth4 = b2.root == B200_EMPTY_ROOT ? 0 : 1;
  if (b2.root < 0) { // empty world or a single leaf: nothing to collapse
    root4 = b2.root;
    return out;
  }
  struct Kid {
    float box[6];
    int32_t ref; // BVH2 reference
  };
  auto area = [](const float *b) {
    const double dx = (double)b[1] - b[0], dy = (double)b[3] - b[2], dz = (double)b[5] - b[4];
    const double a = 2 * (dx * dy + dy * dz + dz * dx);
    return a == a ? a : 0.0;
  };
  struct Job {
    int32_t node2, slot4;
    uint32_t depth;
  };
  std::vector<Job> todo;
  out.emplace_back();
  todo.push_back({b2.root, 0, 1});
  while (!todo.empty()) {
    const Job j = todo.back();
    todo.pop_back();
    max_depth4 = std::max(max_depth4, j.depth + 1);
    std::vector<Kid> kids;
    auto adopt = [&](int32_t n2) {
      const BvhNode &nd = b2.nodes[n2];
      Kid l, r;
      std::memcpy(l.box, nd.lbox, sizeof l.box);
      std::memcpy(r.box, nd.rbox, sizeof r.box);
      l.ref = nd.left, r.ref = nd.right;
      kids.push_back(l);
      kids.push_back(r);
    };
    adopt(j.node2);
    while (kids.size() < 4) {
      int best = -1;
      double best_area = -1;
      for (size_t k = 0; k < kids.size(); k++)
        if (kids[k].ref >= 0 && area(kids[k].box) > best_area) best_area = area(kids[k].box), best = (int)k;
      if (best < 0) break;
      const int32_t n2 = kids[best].ref;
      kids.erase(kids.begin() + best);
      adopt(n2);
    }
    BvhNode4 nd;
    std::memset(&nd, 0, sizeof nd);
    for (int k = 0; k < 4; k++) {
      if (k < (int)kids.size()) {
        nd.x0[k] = kids[k].box[0], nd.x1[k] = kids[k].box[1], nd.y0[k] = kids[k].box[2], nd.y1[k] = kids[k].box[3], nd.z0[k] = kids[k].box[4], nd.z1[k] = kids[k].box[5];
        if (kids[k].ref >= 0) {
          nd.child[k] = (int32_t)out.size();
          out.emplace_back();
          todo.push_back({kids[k].ref, nd.child[k], j.depth + 1});
        } else {
          nd.child[k] = kids[k].ref;
        }
      } else {
        nd.child[k] = B200_EMPTY_ROOT;
      }
    }
    out[j.slot4] = nd;
  }
  root4 = 0;
  return out;
}

} // namespace b200
