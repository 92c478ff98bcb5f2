// scene_dev.cuh — how a scene lives in HBM.
//
// Two views of the same scene are uploaded once (b200rt_upload_scene):
//   (1) the graph of include/b200rt.h verbatim (nodes, list children, geometry rows, materials, textures);
//       the mirror traversal (mirror.cuh) walks it node for node like the reference does;
//   (2) the compacted form the wavefront kernels use: every leaf of the graph becomes a 16-byte Entry
//       {kind | transform chain, geometry row, primitive id, reference visiting order}; entries are stored
//       in the leaf order of our own SAH BVH2 (128-byte nodes that hold both children's boxes), and
//       translate / rotate_* ancestors become a short per-entry chain of ray transforms, applied with the
//       reference's own operations so that t, p and normal come out bit-identical.
// Everything is read-only after upload and small enough to stay L2-resident (DESIGN.md "Data layout").
#pragma once
#include "../../include/b200rt.h"
#include "vec.cuh"

namespace b200 {

enum EntryKind { E_SPHERE = 0, E_SPHERE_MOVING = 1, E_QUAD = 2, E_TRI = 3, E_VOLUME = 4 };

struct __align__(16) Entry {
  uint32_t kind_chain; // kind in bits 0..3, transform chain index in bits 4..31
  int32_t row;         // row in spheres / quads / tris, or index into vol_inst
  int32_t prim;        // primitive id reported by b200rt_trace (node.c), -1 inside a volume boundary
  uint32_t order;      // position in the reference's left-to-right depth-first visiting order (tie-breaks)
};
HD int entry_kind(const Entry &e) { return (int)(e.kind_chain & 15u); }
HD int entry_chain(const Entry &e) { return (int)(e.kind_chain >> 4); }

// One ray transform of a chain: translate (src/hittable.h:75-82) or rotate_x/y/z (:125-149, :192-216, :259-284).
struct __align__(16) ChainOp {
  int32_t kind; // B200RT_NODE_TRANSLATE / ROTATE_X / ROTATE_Y / ROTATE_Z
  int32_t pad;
  double f0, f1, f2; // translate: offset; rotate: f0 = sin, f1 = cos (host-libm values of the reference ctor)
};
struct Chain {
  int32_t first, count; // into chain_ops, outermost transform first
};

// A volumne leaf as visited by the reference: one-object bvh_nodes visit their object twice
// (src/bvh_node.h:34-36), each visit with its own free-flight draw (src/volumne.h:36); `visit` numbers them.
struct __align__(16) VolInst {
  int32_t volume; // row in volumes
  int32_t visit;
  int32_t bfirst, bcount; // boundary entries (closest boundary hit = min t over them)
};

// BVH2 node, 64 bytes = four 128-bit loads: both children's boxes as binary32 planes rounded OUTWARD from the
// (already padded) binary64 boxes.  The box test is a conservative filter only (traverse.cuh); every accepted
// leaf is then tested in binary64, so the fp32 planes never influence a returned t or id.
struct __align__(16) BvhNode {
  float lbox[6]; // left child  {x0,x1,y0,y1,z0,z1}
  float rbox[6]; // right child
  int32_t left, right; // >= 0: inner node index; < 0: leaf, ~ref = first_entry << 3 | (count - 1)
  int32_t pad[2];
};
static_assert(sizeof(BvhNode) == 64, "BvhNode must be half a 128-byte line");

// BVH4 node, 128 bytes = one cache line = eight 128-bit loads: up to four children, planes stored per coordinate so that
// child k's box is lane k of six float4s.  Made by collapsing the BVH2 (bvh_build.h: collapse4); same child references,
// an absent child is B200_EMPTY_ROOT.  Halves the number of DEPENDENT node fetches per ray, which is what a ray pays for
// in trees that do not fit L1 (Sponza-sized scenes: long-scoreboard stalls dominate the BVH2 walk).
struct __align__(16) BvhNode4 {
  float x0[4], x1[4], y0[4], y1[4], z0[4], z1[4];
  int32_t child[4];
  int32_t pad[4];
};
static_assert(sizeof(BvhNode4) == 128, "BvhNode4 must be one 128-byte line");
static_assert(sizeof(Entry) == 16 && sizeof(VolInst) == 16 && sizeof(ChainOp) == 32, "layout");

#define B200_EMPTY_ROOT (-2147483647 - 1)

struct DevScene {
  // (1) graph, verbatim
  const b200rt_node *nodes;
  const uint32_t *list_children;
  uint32_t root;
  const double *spheres;
  const int32_t *sphere_mat;
  const double *quads;
  const int32_t *quad_mat;
  const double *tris;
  const float4 *tri32; // non-null: the nine vertex coordinates of every triangle as binary32 (exactly representable), three float4 per row
  const int32_t *tri_mat;
  const double *tri_uv; // optional: per triangle {u0 v0 u1 v1 u2 v2} (b200rt_scene_desc::tri_uv)
  const b200rt_volume *volumes;
  const b200rt_material *materials;
  const b200rt_texture *textures;
  const b200rt_image *images;
  const uint8_t *image_bytes;
  const double *perlin_vec;
  const int32_t *perlin_perm;
  const b200rt_value_noise *value_noise;
  const float *value_lattice;
  int32_t light_kind, light_index, background_tex;
  // (2) compacted
  const BvhNode *bvh;
  const BvhNode4 *bvh4; // non-null: traverse this (bvh_root indexes it) instead of bvh
  int32_t bvh_root; // child-reference encoding; B200_EMPTY_ROOT for an empty world
  uint32_t n_bvh_nodes;
  const Entry *entries;  // BVH leaf order
  const uint8_t *entry_queue; // per entry: the shade queue of its material (b200rt_material_kind), so that a finished ray needs ONE load, not entry -> row -> material
  const Entry *bentries; // volume boundaries
  const VolInst *vol_inst;
  const ChainOp *chain_ops;
  const Chain *chains;
  uint32_t n_entries;
  int32_t has_volumes;
  const DevScene *self; // this structure again, in global memory (for out-of-line device functions)
  double bound[3]; // max |finite box plane| per axis over the whole BVH: sizes the fp32 error margin of the box test
};

// What extend leaves behind for shade: enough to rebuild the full hit_record deterministically.
struct __align__(16) HitRef {
  double t;
  int32_t entry; // index into entries, -1 = miss, -2 = dead slot (no ray was traced)
  int32_t queue; // which shade queue the path goes to: b200rt_material_kind of the hit, B200_Q_MISS, or -1 for none
};

struct Rec { // src/hittable.h:7-30
  V p, normal;
  double t, u, v;
  bool front_face;
  int mat;
  int prim;
};

} // namespace b200
