// bvh_build.h — host builder of the compacted BVH2 that stands in for the reference's bvh_node tree.
//
// The reference splits at the median of the objects sorted by bounding-box x-min, always on axis 0
// (src/bvh_node.h:20-31, 42-44), which costs ~75 node visits per ray on a field of spheres.
// Here: binned surface-area heuristic over all three axes, leaves of up to `max_leaf` entries, nodes that
// carry both children's boxes so one 128-byte fetch decides both sides.  Which tree is traversed does not
// change the answer: ties are resolved by Entry::order (flatten.h).
#pragma once
#include <chrono>
#include <numeric>

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
  BvhBuilder(const std::vector<Box> &boxes, int max_leaf) : boxes_(boxes), max_leaf_(std::max(1, std::min(8, max_leaf))) {}

  BuiltBvh run() {
    auto t0 = std::chrono::steady_clock::now();
    const size_t n = boxes_.size();
    out_.perm.resize(n);
    std::iota(out_.perm.begin(), out_.perm.end(), 0u);
    cent_.resize(n * 3);
    for (size_t i = 0; i < n; i++)
      for (int k = 0; k < 3; k++) {
        double c = 0.5 * (boxes_[i].lo[k] + boxes_[i].hi[k]);
        cent_[3 * i + k] = std::isfinite(c) ? c : 0.0;
      }
    if (n > 0) {
      Box rb;
      out_.root = build(0, (uint32_t)n, 1, rb);
      root_area_ = rb.area();
      if (root_area_ > 0) out_.sah_cost = cost_acc_ / root_area_;
    }
    out_.build_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    return std::move(out_);
  }

private:
  static constexpr int kBins = 16;
  const std::vector<Box> &boxes_;
  int max_leaf_;
  std::vector<double> cent_;
  BuiltBvh out_;
  double cost_acc_ = 0, root_area_ = 0;

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
  void set_box(float dst[6], const Box &b) {
    for (int k = 0; k < 3; k++) {
      dst[2 * k] = round_down(b.lo[k]);
      dst[2 * k + 1] = round_up(b.hi[k]);
      for (int e = 0; e < 2; e++) {
        const double a = std::fabs((double)dst[2 * k + e]);
        if (std::isfinite(a) && a > out_.bound[k]) out_.bound[k] = a;
      }
    }
  }

  int32_t make_leaf(uint32_t begin, uint32_t end, uint32_t depth, const Box &b) {
    out_.n_leaves++;
    out_.max_depth = std::max(out_.max_depth, depth);
    cost_acc_ += b.area() * (end - begin);
    return ~(int32_t)((begin << 3) | (end - begin - 1));
  }

  int32_t build(uint32_t begin, uint32_t end, uint32_t depth, Box &bounds) {
    const uint32_t count = end - begin;
    bounds.reset();
    Box cb;
    cb.reset();
    for (uint32_t i = begin; i < end; i++) {
      const uint32_t id = out_.perm[i];
      bounds.add(boxes_[id]);
      cb.add(&cent_[3 * id]);
    }
    if (count == 1) return make_leaf(begin, end, depth, bounds);

    // best binned SAH split
    double best_cost = INFINITY;
    int best_axis = -1, best_plane = -1;
    if (depth < 40) {
      for (int axis = 0; axis < 3; axis++) {
        const double lo = cb.lo[axis], ext = cb.hi[axis] - cb.lo[axis];
        if (!(ext > 0) || !std::isfinite(ext)) continue;
        Box bin_box[kBins];
        uint32_t bin_cnt[kBins] = {0};
        for (auto &b : bin_box) b.reset();
        const double scale = kBins / ext;
        for (uint32_t i = begin; i < end; i++) {
          const uint32_t id = out_.perm[i];
          int b = (int)((cent_[3 * id + axis] - lo) * scale);
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
    double parent_area = bounds.area();
    if (!std::isfinite(parent_area)) parent_area = 1e300;
    if (count <= (uint32_t)max_leaf_) {
      // leaf cost: count intersections; split cost: one node (two boxes, ~1.2 intersections' worth) + children
      const double leaf_cost = (double)count * parent_area;
      const double split_cost = best_axis < 0 ? INFINITY : 1.2 * parent_area + best_cost;
      if (leaf_cost <= split_cost) return make_leaf(begin, end, depth, bounds);
    }
    uint32_t mid;
    if (best_axis >= 0) {
      const double lo = cb.lo[best_axis], scale = kBins / (cb.hi[best_axis] - cb.lo[best_axis]);
      auto it = std::partition(out_.perm.begin() + begin, out_.perm.begin() + end, [&](uint32_t id) {
        int b = (int)((cent_[3 * id + best_axis] - lo) * scale);
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
    const uint32_t me = (uint32_t)out_.nodes.size();
    out_.nodes.emplace_back();
    std::memset(&out_.nodes[me], 0, sizeof(BvhNode));
    cost_acc_ += 1.2 * parent_area;
    Box lb, rb;
    const int32_t l = build(begin, mid, depth + 1, lb);
    const int32_t r = build(mid, end, depth + 1, rb);
    BvhNode &nd = out_.nodes[me];
    nd.left = l;
    nd.right = r;
    set_box(nd.lbox, lb);
    set_box(nd.rbox, rb);
    return (int32_t)me;
  }
};

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
