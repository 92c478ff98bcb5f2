// flatten.h — host side of b200rt_upload_scene: graph -> leaf entries with transform chains and world boxes.
//
// Replaces the part of the bvh_node constructor (src/bvh_node.h:18-47) that fixes the ORDER in which
// the reference visits leaves: the walk below is the reference's own left-to-right depth-first visit
// (bvh_node::hit :49-59, hittable_list::hit src/hittable_list.h:20-31), and each leaf's position in it
// becomes Entry::order.  Among hits with equal t the reference returns the one it visited last
// (closed intervals), so (min t, max order) identifies its answer whatever tree we traverse.
#pragma once
#include <algorithm>
#include <cmath>
#include <cstring>
#include <map>
#include <stdexcept>
#include <string>
#include <vector>

#include "scene_dev.cuh"

namespace b200 {

struct Box {
  double lo[3], hi[3];
  void reset() {
    for (int k = 0; k < 3; k++) lo[k] = INFINITY, hi[k] = -INFINITY;
  }
  void add(const double p[3]) {
    for (int k = 0; k < 3; k++) lo[k] = std::min(lo[k], p[k]), hi[k] = std::max(hi[k], p[k]);
  }
  void add(const Box &b) {
    for (int k = 0; k < 3; k++) lo[k] = std::min(lo[k], b.lo[k]), hi[k] = std::max(hi[k], b.hi[k]);
  }
  double area() const {
    const double dx = hi[0] - lo[0], dy = hi[1] - lo[1], dz = hi[2] - lo[2];
    if (!(dx >= 0 && dy >= 0 && dz >= 0)) return 0;
    return 2 * (dx * dy + dy * dz + dz * dx);
  }
};

struct FlatScene {
  std::vector<Entry> entries;  // reference visiting order (BVH build permutes them later)
  std::vector<Box> boxes;      // world boxes, same order, padded
  std::vector<Entry> bentries; // volume boundary leaves
  std::vector<VolInst> vol_inst;
  std::vector<ChainOp> chain_ops;
  std::vector<Chain> chains; // chains[0] = identity
  bool has_volumes = false;
  int graph_depth = 0; // deepest nesting of the reference graph (sizes the device stack of the mirror traversal)
};

class Flattener {
public:
  explicit Flattener(const b200rt_scene_desc &d) : d_(d) {
    out.chains.push_back(Chain{0, 0});
    chain_defs_.push_back({});
  }
  FlatScene out;

  void run() {
    if (d_.n_nodes == 0) return;
    if (d_.root >= d_.n_nodes) throw std::runtime_error("root out of range");
    walk((int)d_.root, 0, 0, false, false, 0);
  }

private:
  const b200rt_scene_desc &d_;
  std::vector<std::vector<ChainOp>> chain_defs_;
  std::map<std::pair<int, int>, int> chain_child_; // (parent chain, node) -> chain
  uint32_t order_ = 0;

  const b200rt_node &node(int i) const {
    if (i < 0 || (uint32_t)i >= d_.n_nodes) throw std::runtime_error("node index out of range");
    return d_.nodes[i];
  }
  int extend_chain(int parent, int node_idx, const ChainOp &op) {
    auto key = std::make_pair(parent, node_idx);
    auto it = chain_child_.find(key);
    if (it != chain_child_.end()) return it->second;
    std::vector<ChainOp> ops = chain_defs_[parent];
    ops.push_back(op);
    Chain c{(int32_t)out.chain_ops.size(), (int32_t)ops.size()};
    out.chain_ops.insert(out.chain_ops.end(), ops.begin(), ops.end());
    out.chains.push_back(c);
    chain_defs_.push_back(ops);
    int id = (int)out.chains.size() - 1;
    chain_child_[key] = id;
    return id;
  }
  // object-space point -> world through chain (inverse of the ray transform), plain arithmetic: boxes are padded
  void to_world(int chain, double p[3]) const {
    const auto &ops = chain_defs_[chain];
    for (int k = (int)ops.size() - 1; k >= 0; k--) {
      const ChainOp &op = ops[k];
      if (op.kind == B200RT_NODE_TRANSLATE) {
        p[0] += op.f0, p[1] += op.f1, p[2] += op.f2;
      } else {
        int i0 = op.kind == B200RT_NODE_ROTATE_X ? 1 : 0, i1 = op.kind == B200RT_NODE_ROTATE_Z ? 1 : 2;
        const double a = p[i0], b = p[i1];
        p[i0] = op.f1 * a + op.f0 * b;
        p[i1] = -op.f0 * a + op.f1 * b;
      }
    }
  }
  static void pad(Box &b) {
    double s = 0;
    for (int k = 0; k < 3; k++) s = std::max(s, std::max(std::fabs(b.lo[k]), std::fabs(b.hi[k])));
    if (!std::isfinite(s)) return;
    const double e = 1e-9 * s + 1e-200;
    for (int k = 0; k < 3; k++) b.lo[k] -= e, b.hi[k] += e;
  }
  Box leaf_box(int kind, int row, int chain) const {
    Box b;
    b.reset();
    if (kind == E_SPHERE || kind == E_SPHERE_MOVING) {
      const double *q = d_.spheres + (size_t)row * B200RT_SPHERE_STRIDE;
      const double r = std::fabs(q[3]);
      const int ncent = kind == E_SPHERE_MOVING ? 2 : 1;
      for (int c = 0; c < ncent; c++) {
        double p[3];
        const double *src = kind == E_SPHERE_MOVING ? q + 5 + 3 * c : q;
        p[0] = src[0], p[1] = src[1], p[2] = src[2];
        to_world(chain, p);
        double lo[3] = {p[0] - r, p[1] - r, p[2] - r}, hi[3] = {p[0] + r, p[1] + r, p[2] + r};
        b.add(lo);
        b.add(hi);
      }
    } else if (kind == E_QUAD) {
      const double *q = d_.quads + (size_t)row * B200RT_QUAD_STRIDE;
      for (int c = 0; c < 4; c++) {
        double p[3];
        for (int k = 0; k < 3; k++) p[k] = q[k] + ((c & 1) ? q[3 + k] : 0.0) + ((c & 2) ? q[6 + k] : 0.0);
        to_world(chain, p);
        b.add(p);
      }
    } else if (kind == E_TRI) {
      const double *q = d_.tris + (size_t)row * B200RT_TRI_STRIDE;
      for (int c = 0; c < 3; c++) {
        double p[3] = {q[3 * c], q[3 * c + 1], q[3 * c + 2]};
        to_world(chain, p);
        b.add(p);
      }
    }
    pad(b);
    return b;
  }

  // second_visit: this subtree is being re-visited through a one-object bvh_node; only volumes (which draw a
  // fresh random number per visit) need a second entry, deterministic leaves would return the same record.
  void walk(int idx, int chain, int visit, bool in_boundary, bool second_visit, int depth) {
    if (depth > 4096) throw std::runtime_error("scene graph too deep (cycle?)");
    out.graph_depth = std::max(out.graph_depth, depth + 1);
    const b200rt_node &n = node(idx);
    switch (n.kind) {
    case B200RT_NODE_BVH: {
      const bool twice = n.a == n.b;
      walk(n.a, chain, twice ? visit * 2 : visit, in_boundary, second_visit, depth + 1);
      walk(n.b, chain, twice ? visit * 2 + 1 : visit, in_boundary, second_visit || twice, depth + 1);
      break;
    }
    case B200RT_NODE_LIST:
      if (n.b < 0 || (uint64_t)n.a + (uint64_t)n.b > d_.n_list_children) throw std::runtime_error("list children out of range");
      for (int k = 0; k < n.b; k++) walk((int)d_.list_children[n.a + k], chain, visit, in_boundary, second_visit, depth + 1);
      break;
    case B200RT_NODE_TRANSLATE:
    case B200RT_NODE_ROTATE_X:
    case B200RT_NODE_ROTATE_Y:
    case B200RT_NODE_ROTATE_Z: {
      ChainOp op;
      std::memset(&op, 0, sizeof op);
      op.kind = n.kind;
      op.f0 = n.f[0], op.f1 = n.f[1], op.f2 = n.f[2];
      walk(n.a, extend_chain(chain, idx, op), visit, in_boundary, second_visit, depth + 1);
      break;
    }
    case B200RT_NODE_SPHERE:
    case B200RT_NODE_QUAD:
    case B200RT_NODE_TRIANGLE: {
      if (second_visit) break;
      int kind;
      uint32_t limit;
      if (n.kind == B200RT_NODE_SPHERE) {
        limit = d_.n_spheres;
        kind = (n.a >= 0 && (uint32_t)n.a < limit && d_.spheres[(size_t)n.a * B200RT_SPHERE_STRIDE + 4] != 0.0) ? E_SPHERE_MOVING : E_SPHERE;
      } else if (n.kind == B200RT_NODE_QUAD) {
        limit = d_.n_quads, kind = E_QUAD;
      } else {
        limit = d_.n_tris, kind = E_TRI;
      }
      if (n.a < 0 || (uint32_t)n.a >= limit) throw std::runtime_error("geometry row out of range");
      Entry e{(uint32_t)kind | ((uint32_t)chain << 4), n.a, in_boundary ? -1 : n.c, order_++};
      if (in_boundary) {
        out.bentries.push_back(e);
        bbox_acc_.add(leaf_box(kind, n.a, chain));
      } else {
        out.entries.push_back(e);
        out.boxes.push_back(leaf_box(kind, n.a, chain));
      }
      break;
    }
    case B200RT_NODE_VOLUME: {
      if (in_boundary) throw std::runtime_error("volumne inside a volumne boundary is unsupported");
      if (n.a < 0 || (uint32_t)n.a >= d_.n_volumes) throw std::runtime_error("volume row out of range");
      const b200rt_volume &vol = d_.volumes[n.a];
      VolInst vi{n.a, visit, (int32_t)out.bentries.size(), 0};
      // boundary leaves carry chains RELATIVE to the volume's own frame (volumne::hit hands them the ray it received)
      bbox_acc_.reset();
      walk(vol.boundary_node, 0, 0, true, false, depth + 1);
      vi.bcount = (int32_t)out.bentries.size() - vi.bfirst;
      out.vol_inst.push_back(vi);
      out.entries.push_back(Entry{(uint32_t)E_VOLUME | ((uint32_t)chain << 4), (int32_t)out.vol_inst.size() - 1, n.c, order_++});
      Box wb;
      wb.reset();
      if (vi.bcount > 0)
        for (int c = 0; c < 8; c++) {
          double p[3] = {(c & 1) ? bbox_acc_.hi[0] : bbox_acc_.lo[0], (c & 2) ? bbox_acc_.hi[1] : bbox_acc_.lo[1], (c & 4) ? bbox_acc_.hi[2] : bbox_acc_.lo[2]};
          to_world(chain, p);
          wb.add(p);
        }
      pad(wb);
      out.boxes.push_back(wb);
      out.has_volumes = true;
      break;
    }
    default: throw std::runtime_error("unknown node kind " + std::to_string(n.kind));
    }
  }
  Box bbox_acc_;
};

} // namespace b200
