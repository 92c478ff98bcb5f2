// api.cu — the C ABI of libb200rt.so (include/b200rt.h): context, scene upload, and the host side of the
// wavefront loop.  No CPU rendering path exists in this library: every entry point that produces rays,
// hits or radiance launches the sm_100a kernels of shade_kernels.cuh / wavefront.cuh.
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <memory>
#include <mutex>

#include "bvh_build.h"
#include "bvh_gpu.cuh"
#include "bvh_ploc.cuh"
#include "output.cuh"
#include "shade_kernels.cuh"

using namespace b200;

namespace {

thread_local std::string t_err;
int fail(int code, const char *fmt, ...) {
  char buf[1024];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof buf, fmt, ap);
  va_end(ap);
  t_err = buf;
  return code;
}
#define CU(x)                                                                                                          \
  do {                                                                                                                 \
    cudaError_t e_ = (x);                                                                                              \
    if (e_ != cudaSuccess) return fail(B200RT_ERR_CUDA, "%s: %s (%s:%d)", #x, cudaGetErrorString(e_), __FILE__, __LINE__); \
  } while (0)

struct DevMem { // owns device allocations of a context
  std::vector<void *> ptrs;
  size_t bytes = 0;
  // optional arena: one contiguous block that the first allocations are carved from, in call order, so that what the
  // traversal touches most (BVH nodes, leaf entries, primitive rows) lies in ONE address range that an L2 access-policy
  // window can cover; what does not fit falls back to cudaMalloc
  char *arena = nullptr;
  size_t arena_size = 0, arena_used = 0;
  ~DevMem() { release(); }
  void release() {
    for (void *p : ptrs) cudaFree(p);
    ptrs.clear();
    bytes = 0;
    arena = nullptr, arena_size = arena_used = 0;
  }
  cudaError_t reserve(size_t n) {
    void *p = nullptr;
    cudaError_t e = cudaMalloc(&p, n);
    if (e == cudaSuccess) ptrs.push_back(p), bytes += n, arena = (char *)p, arena_size = n, arena_used = 0;
    return e;
  }
  cudaError_t alloc(void **p, size_t n) {
    if (n == 0) n = 16;
    const size_t padded = (n + 255) & ~(size_t)255;
    if (arena && arena_used + padded <= arena_size) {
      *p = arena + arena_used;
      arena_used += padded;
      return cudaSuccess;
    }
    cudaError_t e = cudaMalloc(p, n);
    if (e == cudaSuccess) ptrs.push_back(*p), bytes += n;
    return e;
  }
};

} // namespace

#define B200_MAX_PIPELINES 4

struct b200rt_ctx {
  int device = 0;
  int sm_count = 148;
  cudaStream_t stream = nullptr;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;
  // options
  int traversal = B200RT_TRAVERSAL_FAST;
  uint32_t wave_paths = 1u << 25; // shared by the pipelines: 2 x 16 Mi paths = 10 GB of wavefront state
  int kill_zero = 1;
  int bvh_leaf = 0;     // 0 = by scene size (profiles/r01_g_tune.txt): 1 primitive per leaf up to 8192 primitives, else 2
  int refill_lanes = 0; // 0 = by BVH size: fixed 32-ray batches for cache-resident trees, refill at 8 idle lanes from 32 Ki nodes
  int extend_blocks = 4;
  int speculative = 0;
  int fuse_miss = 1;
  int bvh_builder = B200RT_BVH_HOST_SAH;
  int bvh_width = 0; // 0 = by scene size, 2, 4
  uint32_t tail_paths = 32768; // k_tail finishes the frame once every sample is generated and at most this many paths live (0 = off)
  int extend_stages = 2; // 1 is within 1 % on RTOW / RTNW / Cornell and 11 % slower on Fox (profiles/r01_h_tune_l1.txt)
  int carveout_pct = -1; // -1 = computed from the kernel's footprint
  // scene
  bool have_scene = false, have_camera = false;
  DevMem scene_mem;
  DevScene S{};
  b200rt_bvh_info bvh_info{};
  double scene_h2d_bytes = 0;
  int graph_depth = 0;
  int bg_mode = 0; // WaveParams::fused_miss for this scene
  bool scene_optins = false; // the scene uses float images, bilinear lookups or triangle texture coordinates: the EXT kernels run
  uint32_t n_materials = 0;
  uint32_t kinds_present = 0; // bit k: some material of the scene has b200rt_material_kind k (queues of absent kinds stay empty)
  // camera
  b200rt_camera cam{};
  // wavefront state: one set per pipeline (B200RT_OPT_PIPELINES; set 1 is only allocated for two pipelines)
  struct WaveSet {
    DevMem mem;
    uint32_t capacity = 0;
    bool mirror = false;
    Pools pools{};
    HitRef *hits = nullptr;
    Rec *recs = nullptr;
    Queues queues{};
    WaveCtl *ctl = nullptr;
    uint32_t *sort_keys[2] = {nullptr, nullptr}, *sort_vals[2] = {nullptr, nullptr}; // B200RT_OPT_RAY_SORT (experiment)
    void *sort_tmp = nullptr;
    size_t sort_tmp_bytes = 0;
    WaveCtl *h_ctl = nullptr; // pinned
    cudaStream_t stream = nullptr; // pipeline stream (two-pipeline mode)
    cudaEvent_t done = nullptr;
    // device-driven wave loop: one CUDA graph = a while node whose body is one wave; rebuilt when anything baked into its
    // kernel nodes (scene pointers, camera, sample range, seed, accumulator, options) changes
    cudaGraph_t graph = nullptr;
    cudaGraphExec_t graph_exec = nullptr;
    std::string graph_key;
    int graph_launches_per_wave = 0;
    void drop_graph() {
      if (graph_exec) cudaGraphExecDestroy(graph_exec);
      if (graph) cudaGraphDestroy(graph);
      graph_exec = nullptr, graph = nullptr;
      graph_key.clear();
    }
  } w[B200_MAX_PIPELINES];
  cudaStream_t cap_stream = nullptr; // stream captures are recorded on (never executes anything)
  int l2_persist = 1;  // B200RT_OPT_L2_PERSIST
  int tri_f32 = 1;     // B200RT_OPT_TRI_F32
  size_t hot_bytes = 0; // prefix of the scene arena that the traversal reads per node / leaf
  size_t l2_window_max = 0, l2_persist_max = 0;
  int ray_sort = 0;    // B200RT_OPT_RAY_SORT
  int octant_sort = 0; // B200RT_OPT_OCTANT_SORT
  int extend_loop = -1; // B200RT_OPT_EXTEND_LOOP: 0 = while-while, 1 = warp-uniform threshold if-if, -1 = by scene (1 on triangle meshes)
  int leaf_lanes = 0;   // B200RT_OPT_LEAF_LANES; 0 = 20 when the scene picked the loop, 8 when the caller did
  bool mesh_scene = false; // a BVH4 whose leaves are almost all triangles (set by b200rt_upload_scene)
  int loop_mode() const { return extend_loop >= 0 ? extend_loop : (mesh_scene ? 1 : 0); }
  int loop_lanes() const { return leaf_lanes > 0 ? leaf_lanes : (extend_loop < 0 ? 20 : 8); }
  int wave_loop = 1; // B200RT_OPT_WAVE_LOOP: 1 = device-driven CUDA-graph while loop, 0 = host-driven loop with read-backs
  cudaEvent_t ev_fork = nullptr;
  int pipelines = 2;        // measured: +4 % on RTOW, +5 % on the Sponza stand-in over one pipeline (profiles/r01_r_pipelines.txt)
  int pipe_extend_ctas = 4; // fewer (leaving registers for the other half's kernels) is slower: extend needs its occupancy
  double *accum = nullptr;  // W*H*3, used by b200rt_render
  size_t accum_elems = 0;
};

namespace {

template <class T> int upload(b200rt_ctx *c, const T *src, size_t count, const T **dst, const char *what) {
  void *p = nullptr;
  if (count && !src) return fail(B200RT_ERR_INVALID, "scene description: %s is NULL but its count is %zu", what, count);
  CU(c->scene_mem.alloc(&p, count * sizeof(T)));
  if (count) CU(cudaMemcpyAsync(p, src, count * sizeof(T), cudaMemcpyHostToDevice, c->stream));
  c->scene_h2d_bytes += (double)(count * sizeof(T));
  *dst = static_cast<const T *>(p);
  return B200RT_OK;
}
#define UP(src, count, dst, what)                                                                                      \
  do {                                                                                                                 \
    int rc_ = upload(c, src, (size_t)(count), dst, what);                                                              \
    if (rc_ != B200RT_OK) return rc_;                                                                                  \
  } while (0)

#ifndef B200_AGGLOMERATIVE_MAX_PRIMS
#define B200_AGGLOMERATIVE_MAX_PRIMS 64 /* scenes up to this size also get the exact agglomerative tree (O(n^3) on the host: < 1 ms at 64, 40 ms at 256) */
#endif
#ifndef B200_WIDE_FROM_PRIMS
#define B200_WIDE_FROM_PRIMS 2048 /* scenes above this many primitives are traversed as a BVH4 (their BVH2 does not fit L1) */
#endif

// Device LBVH (bvh_gpu.cuh): leaves the finished nodes and the entries in leaf order in scene memory.
int build_bvh_device(b200rt_ctx *c, const FlatScene &fs, DevScene &S, b200rt_bvh_info &info) {
  const uint32_t n = (uint32_t)fs.boxes.size();
  if (n >= (1u << 28)) return fail(B200RT_ERR_UNSUPPORTED, "device BVH build: %u primitives exceed the leaf reference range", n);
  DevMem tmp;
  cudaStream_t s = c->stream;
  Box *d_boxes = nullptr;
  Entry *d_in = nullptr, *d_out = nullptr;
  uint32_t *codes[2] = {nullptr, nullptr}, *ids[2] = {nullptr, nullptr};
  int2 *children = nullptr;
  int *parent_inner = nullptr, *parent_leaf = nullptr;
  unsigned int *arrived = nullptr;
  Box32 *node_box = nullptr;
  LbvhStats *st = nullptr;
  BvhNode *nodes = nullptr;
  const size_t ni = n > 1 ? n - 1 : 0;
  CU(tmp.alloc((void **)&d_boxes, n * sizeof(Box)));
  CU(tmp.alloc((void **)&d_in, n * sizeof(Entry)));
  for (int k = 0; k < 2; k++) {
    CU(tmp.alloc((void **)&codes[k], n * 4));
    CU(tmp.alloc((void **)&ids[k], n * 4));
  }
  CU(tmp.alloc((void **)&children, ni * sizeof(int2)));
  CU(tmp.alloc((void **)&parent_inner, ni * 4));
  CU(tmp.alloc((void **)&parent_leaf, n * 4));
  CU(tmp.alloc((void **)&arrived, ni * 4));
  CU(tmp.alloc((void **)&node_box, ni * sizeof(Box32)));
  CU(tmp.alloc((void **)&st, sizeof(LbvhStats)));
  CU(c->scene_mem.alloc((void **)&nodes, ni * sizeof(BvhNode)));
  CU(c->scene_mem.alloc((void **)&d_out, n * sizeof(Entry)));
  size_t cub_bytes = 0;
  CU(cub::DeviceRadixSort::SortPairs(nullptr, cub_bytes, codes[0], codes[1], ids[0], ids[1], (int)n, 0, 30, s));
  void *cub_tmp = nullptr;
  CU(tmp.alloc(&cub_tmp, cub_bytes));
  CU(cudaMemcpyAsync(d_boxes, fs.boxes.data(), n * sizeof(Box), cudaMemcpyHostToDevice, s));
  CU(cudaMemcpyAsync(d_in, fs.entries.data(), n * sizeof(Entry), cudaMemcpyHostToDevice, s));
  c->scene_h2d_bytes += (double)n * (sizeof(Box) + sizeof(Entry));
  CU(cudaEventRecord(c->ev0, s));
  const int g = (int)((n + 255) / 256);
  k_lbvh_init<<<1, 1, 0, s>>>(st);
  k_lbvh_bounds<<<std::min(g, c->sm_count * 8), 256, 0, s>>>(d_boxes, n, st);
  k_lbvh_morton<<<g, 256, 0, s>>>(d_boxes, n, st, codes[0], ids[0]);
  CU(cub::DeviceRadixSort::SortPairs(cub_tmp, cub_bytes, codes[0], codes[1], ids[0], ids[1], (int)n, 0, 30, s));
  CU(cudaMemsetAsync(parent_leaf, 0xFF, n * 4, s));
  CU(cudaMemsetAsync(arrived, 0, ni * 4, s));
  if (n > 1) k_lbvh_hierarchy<<<(int)((ni + 255) / 256), 256, 0, s>>>(codes[1], (int)n, children, parent_inner, parent_leaf);
  k_lbvh_refit<<<g, 256, 0, s>>>(d_boxes, ids[1], (int)n, children, parent_inner, parent_leaf, arrived, node_box, nodes, st);
  k_lbvh_gather<<<g, 256, 0, s>>>(d_in, ids[1], n, d_out);
  CU(cudaEventRecord(c->ev1, s));
  CU(cudaGetLastError());
  LbvhStats h;
  CU(cudaMemcpyAsync(&h, st, sizeof h, cudaMemcpyDeviceToHost, s));
  CU(cudaStreamSynchronize(s));
  float ms = 0;
  CU(cudaEventElapsedTime(&ms, c->ev0, c->ev1));
  S.bvh = nodes;
  S.entries = d_out;
  S.bvh_root = n > 1 ? 0 : ~0; // inner node 0 is the root of a Karras tree; a single primitive is a leaf reference
  S.n_bvh_nodes = (uint32_t)ni;
  S.n_entries = n;
  for (int k = 0; k < 3; k++) {
    float f;
    std::memcpy(&f, &h.plane_max[k], 4);
    S.bound[k] = (double)f;
  }
  info.n_nodes = (uint32_t)ni;
  info.n_leaves = n;
  info.n_prims = n;
  info.max_depth = h.max_depth;
  info.build_ms = ms;
  info.sah_cost = h.root_area > 0 ? h.sah_acc / h.root_area : 0.0;
  if (h.max_depth > B200_STACK) return fail(B200RT_ERR_UNSUPPORTED, "device BVH build: depth %u exceeds the traversal stack", h.max_depth);
  return B200RT_OK;
}

// Device PLOC (bvh_ploc.cuh) + host finishing pass (bvh_build.h:finish_tree): the BVH2 topology is clustered on the device, comes back
// (64 bytes per node), gets its leaf pairs, depth-first layout and, in b200rt_upload_scene, the BVH4 collapse on the host.  `out.perm`
// indexes the flattener's entry order like the host builder's.  Returns B200RT_ERR_UNSUPPORTED if the clustering does not converge in
// a sane number of rounds (the caller then falls back to the host builder).
int build_bvh_ploc(b200rt_ctx *c, const FlatScene &fs, int max_leaf, BuiltBvh &out) {
  const uint32_t n = (uint32_t)fs.boxes.size();
  if (n >= (1u << 28)) return fail(B200RT_ERR_UNSUPPORTED, "device BVH build: %u primitives exceed the leaf reference range", n);
  DevMem tmp;
  cudaStream_t s = c->stream;
  Box *d_boxes = nullptr;
  uint32_t *codes[2] = {nullptr, nullptr}, *ids[2] = {nullptr, nullptr};
  PlocCluster *cl[2] = {nullptr, nullptr};
  int *nn = nullptr;
  unsigned long long *flags = nullptr, *scan = nullptr;
  unsigned int *totals = nullptr;
  LbvhStats *st = nullptr;
  BvhNode *nodes = nullptr;
  const size_t ni = n > 1 ? n - 1 : 0;
  CU(tmp.alloc((void **)&d_boxes, n * sizeof(Box)));
  for (int k = 0; k < 2; k++) {
    CU(tmp.alloc((void **)&codes[k], n * 4));
    CU(tmp.alloc((void **)&ids[k], n * 4));
    CU(tmp.alloc((void **)&cl[k], n * sizeof(PlocCluster)));
  }
  CU(tmp.alloc((void **)&nn, n * 4));
  CU(tmp.alloc((void **)&flags, n * 8));
  CU(tmp.alloc((void **)&scan, n * 8));
  CU(tmp.alloc((void **)&totals, 8));
  CU(tmp.alloc((void **)&st, sizeof(LbvhStats)));
  CU(tmp.alloc((void **)&nodes, ni * sizeof(BvhNode)));
  size_t sort_bytes = 0, scan_bytes = 0;
  CU(cub::DeviceRadixSort::SortPairs(nullptr, sort_bytes, codes[0], codes[1], ids[0], ids[1], (int)n, 0, 30, s));
  CU(cub::DeviceScan::ExclusiveSum(nullptr, scan_bytes, flags, scan, (int)n, s));
  void *cub_tmp = nullptr;
  CU(tmp.alloc(&cub_tmp, std::max(sort_bytes, scan_bytes)));
  CU(cudaMemcpyAsync(d_boxes, fs.boxes.data(), n * sizeof(Box), cudaMemcpyHostToDevice, s));
  c->scene_h2d_bytes += (double)n * sizeof(Box);
  CU(cudaEventRecord(c->ev0, s));
  const int g = (int)((n + 255) / 256);
  k_lbvh_init<<<1, 1, 0, s>>>(st);
  k_lbvh_bounds<<<std::min(g, c->sm_count * 8), 256, 0, s>>>(d_boxes, n, st);
  k_lbvh_morton<<<g, 256, 0, s>>>(d_boxes, n, st, codes[0], ids[0]);
  CU(cub::DeviceRadixSort::SortPairs(cub_tmp, sort_bytes, codes[0], codes[1], ids[0], ids[1], (int)n, 0, 30, s));
  k_ploc_init<<<g, 256, 0, s>>>(d_boxes, ids[1], n, cl[0], st);
  uint32_t m = n, node_base = 0;
  int cur = 0, rounds = 0;
  const char *env_radius = std::getenv("B200RT_PLOC_RADIUS"); // experiments only
  const int radius = env_radius && std::atoi(env_radius) > 0 ? std::atoi(env_radius) : B200_PLOC_RADIUS;
  while (m > 1) {
    if (++rounds > 4096) return fail(B200RT_ERR_UNSUPPORTED, "device PLOC build: %u clusters left after %d rounds", m, rounds - 1);
    const int gm = (int)((m + 255) / 256);
    k_ploc_nearest<<<gm, 256, 0, s>>>(cl[cur], (int)m, radius, nn);
    k_ploc_flag<<<gm, 256, 0, s>>>(nn, (int)m, flags);
    CU(cub::DeviceScan::ExclusiveSum(cub_tmp, scan_bytes, flags, scan, (int)m, s));
    k_ploc_merge<<<gm, 256, 0, s>>>(cl[cur], nn, flags, scan, (int)m, node_base, cl[cur ^ 1], nodes, totals);
    unsigned int h[2] = {0, 0};
    CU(cudaMemcpyAsync(h, totals, 8, cudaMemcpyDeviceToHost, s));
    CU(cudaStreamSynchronize(s));
    if (h[1] == 0 || h[0] + h[1] != m) return fail(B200RT_ERR_CUDA, "device PLOC build: round %d kept %u and merged %u of %u clusters", rounds, h[0], h[1], m);
    m = h[0], node_base += h[1], cur ^= 1;
  }
  CU(cudaEventRecord(c->ev1, s));
  CU(cudaGetLastError());
  if (node_base != ni) return fail(B200RT_ERR_CUDA, "device PLOC build: %u nodes for %u primitives", node_base, n);
  std::vector<BvhNode> topo(ni);
  std::vector<uint32_t> sorted(n);
  LbvhStats h;
  CU(cudaMemcpyAsync(topo.data(), nodes, ni * sizeof(BvhNode), cudaMemcpyDeviceToHost, s));
  CU(cudaMemcpyAsync(sorted.data(), ids[1], n * 4, cudaMemcpyDeviceToHost, s));
  CU(cudaMemcpyAsync(&h, st, sizeof h, cudaMemcpyDeviceToHost, s));
  CU(cudaStreamSynchronize(s));
  float ms = 0;
  CU(cudaEventElapsedTime(&ms, c->ev0, c->ev1));
  double bound[3];
  for (int k = 0; k < 3; k++) {
    float f;
    std::memcpy(&f, &h.plane_max[k], 4);
    bound[k] = (double)f;
  }
  out = finish_tree(topo, n, max_leaf, bound);
  for (uint32_t &p : out.perm) p = sorted[p]; // leaf order -> Morton position -> flattener's entry order
  out.build_ms += ms;                         // device clustering + host finishing pass (the copies in between are not in either)
  return B200RT_OK;
}

int validate(const b200rt_scene_desc *d) {
  if (!d) return fail(B200RT_ERR_INVALID, "scene description is NULL");
  if (d->abi_version != B200RT_ABI_VERSION) return fail(B200RT_ERR_INVALID, "scene description has ABI version %u, library has %d", d->abi_version, B200RT_ABI_VERSION);
  if (d->n_nodes && d->root >= d->n_nodes) return fail(B200RT_ERR_INVALID, "root %u out of range (%u nodes)", d->root, d->n_nodes);
  for (uint32_t i = 0; i < d->n_volumes; i++) {
    if (d->volumes[i].material < 0 || (uint32_t)d->volumes[i].material >= d->n_materials) return fail(B200RT_ERR_INVALID, "volume %u: material out of range", i);
    if (d->volumes[i].boundary_node < 0 || (uint32_t)d->volumes[i].boundary_node >= d->n_nodes) return fail(B200RT_ERR_INVALID, "volume %u: boundary node out of range", i);
  }
  for (uint32_t i = 0; i < d->n_materials; i++) {
    const b200rt_material &m = d->materials[i];
    if (m.kind < 0 || m.kind > B200RT_MAT_DIFFUSE_LIGHT) return fail(B200RT_ERR_UNSUPPORTED, "material %u: unknown kind %d", i, m.kind);
    if (m.tex < 0 || (uint32_t)m.tex >= d->n_textures) return fail(B200RT_ERR_INVALID, "material %u: texture out of range", i);
  }
  for (uint32_t i = 0; i < d->n_textures; i++) {
    const b200rt_texture &t = d->textures[i];
    if (t.kind < 0 || t.kind > B200RT_TEX_PICTURE_LINEAR) return fail(B200RT_ERR_UNSUPPORTED, "texture %u: unknown kind %d", i, t.kind);
    if ((t.kind == B200RT_TEX_PICTURE || t.kind == B200RT_TEX_PICTURE_LINEAR) && (t.aux < 0 || (uint32_t)t.aux >= d->n_images)) return fail(B200RT_ERR_INVALID, "texture %u: image out of range", i);
    if (t.kind == B200RT_TEX_PERLIN && (t.aux < 0 || (uint32_t)t.aux >= d->n_perlin)) return fail(B200RT_ERR_INVALID, "texture %u: perlin table out of range", i);
    if (t.kind == B200RT_TEX_VALUE && (t.aux < 0 || (uint32_t)t.aux >= d->n_value_noise)) return fail(B200RT_ERR_INVALID, "texture %u: value-noise table out of range", i);
  }
  for (uint32_t i = 0; i < d->n_images; i++) {
    const b200rt_image &im = d->images[i];
    if (im.has_data && (im.width <= 0 || im.height <= 0 || (im.bytes_per_pixel != 3 && im.bytes_per_pixel != 4 && im.bytes_per_pixel != 12 && im.bytes_per_pixel != 16) ||
                        (im.bytes_per_pixel >= 12 && im.offset % 4 != 0) ||
                        im.offset + (uint64_t)im.width * im.height * im.bytes_per_pixel > d->image_bytes_size))
      return fail(B200RT_ERR_INVALID, "image %u: bad geometry or bytes out of range", i);
  }
  if (d->background_tex >= (int32_t)d->n_textures) return fail(B200RT_ERR_INVALID, "background texture out of range");
  if (d->light_kind == B200RT_LIGHT_QUAD && (d->light_index < 0 || (uint32_t)d->light_index >= d->n_quads)) return fail(B200RT_ERR_INVALID, "light quad out of range");
  if (d->light_kind == B200RT_LIGHT_SPHERE && (d->light_index < 0 || (uint32_t)d->light_index >= d->n_spheres)) return fail(B200RT_ERR_INVALID, "light sphere out of range");
  if (d->light_kind < 0 || d->light_kind > B200RT_LIGHT_DEFAULT) return fail(B200RT_ERR_INVALID, "unknown light kind %d", d->light_kind);
  return B200RT_OK;
}

} // namespace

// ================================================================================================ C ABI
extern "C" {

const char *b200rt_last_error(void) { return t_err.c_str(); }
int b200rt_abi_version(void) { return B200RT_ABI_VERSION; }

// sizes of the ABI structures as compiled; the ctypes mirror checks itself against these
size_t b200rt_abi_sizeof(int which) {
  switch (which) {
  case 0: return sizeof(b200rt_scene_desc);
  case 1: return sizeof(b200rt_camera);
  case 2: return sizeof(b200rt_stats);
  case 3: return sizeof(b200rt_node);
  case 4: return sizeof(b200rt_material);
  case 5: return sizeof(b200rt_texture);
  case 6: return sizeof(b200rt_image);
  case 7: return sizeof(b200rt_volume);
  case 8: return sizeof(b200rt_value_noise);
  case 9: return sizeof(b200rt_bvh_info);
  }
  return 0;
}

int b200rt_create(int device, b200rt_ctx **out) {
  if (!out) return fail(B200RT_ERR_INVALID, "b200rt_create: out is NULL");
  *out = nullptr;
  int n = 0;
  CU(cudaGetDeviceCount(&n));
  if (device < 0 || device >= n) return fail(B200RT_ERR_INVALID, "b200rt_create: device %d out of range (%d CUDA devices)", device, n);
  CU(cudaSetDevice(device));
  cudaDeviceProp prop;
  CU(cudaGetDeviceProperties(&prop, device));
  if (prop.major < 10) return fail(B200RT_ERR_UNSUPPORTED, "b200rt_create: device %d is sm_%d%d; this library carries sm_100a code only", device, prop.major, prop.minor);
  std::unique_ptr<b200rt_ctx> c(new b200rt_ctx);
  c->device = device;
  c->sm_count = prop.multiProcessorCount;
  c->l2_window_max = (size_t)std::max(0, prop.accessPolicyMaxWindowSize);
  c->l2_persist_max = (size_t)std::max(0, prop.persistingL2CacheMaxSize);
  CU(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
  CU(cudaEventCreate(&c->ev0));
  CU(cudaEventCreate(&c->ev1));
  CU(cudaEventCreateWithFlags(&c->ev_fork, cudaEventDisableTiming));
  CU(cudaStreamCreateWithFlags(&c->cap_stream, cudaStreamNonBlocking));
  if (const char *e = getenv("B200RT_WAVE_LOOP")) c->wave_loop = atoi(e) != 0; // debugging aid: default of B200RT_OPT_WAVE_LOOP
  for (auto &w : c->w) {
    CU(cudaMallocHost(&w.h_ctl, sizeof(WaveCtl)));
    CU(cudaStreamCreateWithFlags(&w.stream, cudaStreamNonBlocking));
    CU(cudaEventCreateWithFlags(&w.done, cudaEventDisableTiming));
  }
  *out = c.release();
  return B200RT_OK;
}

int b200rt_destroy(b200rt_ctx *c) {
  if (!c) return B200RT_OK;
  cudaSetDevice(c->device);
  if (c->stream) cudaStreamSynchronize(c->stream);
  c->scene_mem.release();
  for (auto &w : c->w) {
    if (w.stream) cudaStreamSynchronize(w.stream);
    w.drop_graph();
    w.mem.release();
    if (w.h_ctl) cudaFreeHost(w.h_ctl);
    if (w.done) cudaEventDestroy(w.done);
    if (w.stream) cudaStreamDestroy(w.stream);
  }
  if (c->ev_fork) cudaEventDestroy(c->ev_fork);
  if (c->cap_stream) cudaStreamDestroy(c->cap_stream);
  if (c->ev0) cudaEventDestroy(c->ev0);
  if (c->ev1) cudaEventDestroy(c->ev1);
  if (c->stream) cudaStreamDestroy(c->stream);
  delete c;
  return B200RT_OK;
}

int b200rt_set_option(b200rt_ctx *c, int option, double value) {
  if (!c) return fail(B200RT_ERR_INVALID, "context is NULL");
  switch (option) {
  case B200RT_OPT_TRAVERSAL:
    if (value != B200RT_TRAVERSAL_FAST && value != B200RT_TRAVERSAL_MIRROR) return fail(B200RT_ERR_INVALID, "traversal must be 0 (compacted BVH) or 1 (mirror)");
    c->traversal = (int)value;
    return B200RT_OK;
  case B200RT_OPT_WAVE_PATHS:
    if (!(value >= 1024 && value <= 268435456.0)) return fail(B200RT_ERR_INVALID, "wave paths must be in [1024, 2^28]");
    c->wave_paths = (uint32_t)value;
    return B200RT_OK;
  case B200RT_OPT_KILL_ZERO_WEIGHT: c->kill_zero = value != 0; return B200RT_OK;
  case B200RT_OPT_BVH_LEAF_SIZE:
    if (!(value >= 0 && value <= 8)) return fail(B200RT_ERR_INVALID, "leaf size must be in [0, 8] (0 = automatic)");
    c->bvh_leaf = (int)value; // takes effect at the next b200rt_upload_scene
    return B200RT_OK;
  case B200RT_OPT_EXTEND_BLOCKS: c->extend_blocks = (int)value; return B200RT_OK;
  case B200RT_OPT_SPECULATIVE_LEAF: c->speculative = value != 0; return B200RT_OK;
  case B200RT_OPT_FUSE_MISS: c->fuse_miss = value != 0; return B200RT_OK;
  case B200RT_OPT_PIPELINES:
    if (!(value >= 1 && value <= B200_MAX_PIPELINES)) return fail(B200RT_ERR_INVALID, "pipelines must be in [1, %d]", B200_MAX_PIPELINES);
    c->pipelines = (int)value;
    return B200RT_OK;
  case B200RT_OPT_PIPE_EXTEND_CTAS:
    if (!(value >= 1 && value <= 4)) return fail(B200RT_ERR_INVALID, "extend CTAs per SM in two-pipeline mode must be in [1, 4]");
    c->pipe_extend_ctas = (int)value;
    return B200RT_OK;
  case B200RT_OPT_BVH_WIDTH:
    if (value != 0 && value != 2 && value != 4) return fail(B200RT_ERR_INVALID, "BVH width must be 0 (automatic), 2 or 4");
    c->bvh_width = (int)value; // takes effect at the next b200rt_upload_scene
    return B200RT_OK;
  case B200RT_OPT_BVH_BUILDER:
    if (value != B200RT_BVH_HOST_SAH && value != B200RT_BVH_DEVICE_LBVH && value != B200RT_BVH_DEVICE_PLOC)
      return fail(B200RT_ERR_INVALID, "BVH builder must be 0 (host SAH), 1 (device LBVH) or 2 (device PLOC)");
    c->bvh_builder = (int)value; // takes effect at the next b200rt_upload_scene
    return B200RT_OK;
  case B200RT_OPT_TAIL_PATHS:
    if (!(value >= 0 && value <= 1048576.0)) return fail(B200RT_ERR_INVALID, "tail paths must be in [0, 2^20]");
    c->tail_paths = (uint32_t)value;
    return B200RT_OK;
  case B200RT_OPT_EXTEND_STAGES:
    if (value != 1 && value != 2) return fail(B200RT_ERR_INVALID, "extend stages must be 1 or 2");
    c->extend_stages = (int)value;
    return B200RT_OK;
  case B200RT_OPT_SMEM_CARVEOUT:
    if (!(value >= -1 && value <= 100)) return fail(B200RT_ERR_INVALID, "carve-out must be -1 (automatic) or a percentage");
    c->carveout_pct = (int)value;
    return B200RT_OK;
  case B200RT_OPT_L2_PERSIST: c->l2_persist = value != 0; return B200RT_OK;
  case B200RT_OPT_TRI_F32: c->tri_f32 = value != 0; return B200RT_OK; // takes effect at the next b200rt_upload_scene // takes effect at the next b200rt_upload_scene
  case B200RT_OPT_RAY_SORT:
    if (!(value >= 0 && value <= 7)) return fail(B200RT_ERR_INVALID, "ray sort mode must be in [0, 7]");
#if !B200_RAY_SORT_EXPERIMENT
    if (value != 0) return fail(B200RT_ERR_UNSUPPORTED, "the ray-reordering experiment is not compiled into this build (-DB200_RAY_SORT_EXPERIMENT=1)");
#endif
    c->ray_sort = (int)value;
    for (auto &w : c->w) w.capacity = 0; // the sort buffers are allocated with the wave pools
    return B200RT_OK;
  case B200RT_OPT_OCTANT_SORT:
#if !B200_OCTANT_SORT_EXPERIMENT
    if (value != 0) return fail(B200RT_ERR_UNSUPPORTED, "the octant-grouping experiment is not compiled into this build (-DB200_OCTANT_SORT_EXPERIMENT=1)");
#endif
    c->octant_sort = value != 0;
    return B200RT_OK;
  case B200RT_OPT_EXTEND_LOOP:
    if (value != 0 && value != 1 && value != -1) return fail(B200RT_ERR_INVALID, "extend loop must be 0 (while-while), 1 (warp-uniform) or -1 (chosen by scene)");
    c->extend_loop = (int)value;
    return B200RT_OK;
  case B200RT_OPT_LEAF_LANES:
    if (!(value >= 0 && value <= 32)) return fail(B200RT_ERR_INVALID, "leaf lanes must be in [1, 32] (0 = default)");
    c->leaf_lanes = (int)value;
    return B200RT_OK;
  case B200RT_OPT_WAVE_LOOP:
    if (value != 0 && value != 1) return fail(B200RT_ERR_INVALID, "wave loop must be 0 (host-driven) or 1 (device-driven CUDA graph)");
    c->wave_loop = (int)value;
    return B200RT_OK;
  case B200RT_OPT_REFILL_LANES:
    if (!(value >= 0 && value <= 32)) return fail(B200RT_ERR_INVALID, "refill lanes must be in [0, 32] (0 = automatic)");
    c->refill_lanes = (int)value;
    return B200RT_OK;
  }
  return fail(B200RT_ERR_INVALID, "unknown option %d", option);
}

int b200rt_upload_scene(b200rt_ctx *c, const b200rt_scene_desc *d) {
  if (!c) return fail(B200RT_ERR_INVALID, "context is NULL");
  int rc = validate(d);
  if (rc != B200RT_OK) return rc;
  CU(cudaSetDevice(c->device));
  CU(cudaStreamSynchronize(c->stream));
  c->have_scene = false;
  c->scene_mem.release();
  c->scene_h2d_bytes = 0;

  // host: reference visiting order + transform chains + world boxes, then our own BVH
  FlatScene fs;
  BuiltBvh bvh;
  std::vector<Entry> ordered;
  bool use_device_builder = false;
  uint32_t n_nodes4 = 0;
  try {
    Flattener fl(*d);
    fl.run();
    fs = std::move(fl.out);
    for (const Entry &e : fs.entries) {
      int mat = -1;
      switch (entry_kind(e)) {
      case E_SPHERE:
      case E_SPHERE_MOVING: mat = d->sphere_mat[e.row]; break;
      case E_QUAD: mat = d->quad_mat[e.row]; break;
      case E_TRI: mat = d->tri_mat[e.row]; break;
      default: mat = 0; break;
      }
      if (mat < 0 || (uint32_t)mat >= d->n_materials)
        return fail(B200RT_ERR_UNSUPPORTED, "primitive %d has no material (the reference would dereference a null pointer)", e.prim);
    }
    if (!(use_device_builder = c->bvh_builder == B200RT_BVH_DEVICE_LBVH && !fs.boxes.empty())) {
      const int max_leaf = c->bvh_leaf ? c->bvh_leaf : (fs.boxes.size() <= 8192 ? 1 : 2);
      bool built = false;
      if (c->bvh_builder == B200RT_BVH_DEVICE_PLOC && !fs.boxes.empty()) {
        const int prc = build_bvh_ploc(c, fs, max_leaf, bvh);
        if (prc != B200RT_OK && prc != B200RT_ERR_UNSUPPORTED) return prc;
        built = prc == B200RT_OK && bvh.max_depth <= B200_STACK; // a degenerate clustering (too deep, no convergence): host builder
      }
      // experiments: B200RT_SAH_PEEL=0 switches the largest-primitive candidate off, B200RT_SAH_SWEEP=1 the exact sweep on small nodes on,
      // B200RT_SAH_AREA=1 the box-area axis
      auto env_is = [](const char *name, int dflt) { const char *v = std::getenv(name); return (v ? std::atoi(v) : dflt) != 0; };
      if (!built) {
        bvh = BvhBuilder(fs.boxes, max_leaf, env_is("B200RT_SAH_PEEL", 1), env_is("B200RT_SAH_SWEEP", 0), env_is("B200RT_SAH_AREA", 0)).run();
        // a few large overlapping primitives (a room and what stands in it): bottom-up merging often finds the cheaper tree
        if (fs.boxes.size() >= 3 && fs.boxes.size() <= B200_AGGLOMERATIVE_MAX_PRIMS && env_is("B200RT_AGGLOMERATIVE", 1)) {
          BuiltBvh alt = build_agglomerative(fs.boxes, max_leaf);
          if (alt.max_depth <= B200_STACK && alt.sah_cost < 0.99 * bvh.sah_cost) {
            alt.build_ms += bvh.build_ms;
            bvh = std::move(alt);
          }
        }
      }
      ordered.resize(fs.entries.size());
      for (size_t k = 0; k < ordered.size(); k++) ordered[k] = fs.entries[bvh.perm[k]];
      // node_step pushes at most one reference per level: a tree deeper than the traversal stack would drop subtrees silently
      if (bvh.max_depth > B200_STACK)
        return fail(B200RT_ERR_UNSUPPORTED, "host BVH build: depth %u exceeds the traversal stack (%d)", bvh.max_depth, B200_STACK);
    }
  } catch (const std::exception &e) {
    return fail(B200RT_ERR_INVALID, "scene description: %s", e.what());
  }

  DevScene S;
  std::memset(&S, 0, sizeof S);
  S.root = d->root;
  // the tree that is traversed: the BVH2, or the BVH4 collapsed from it
  std::vector<BvhNode4> wide;
  if (!use_device_builder) {
    const int width = c->bvh_width ? c->bvh_width : (ordered.size() > B200_WIDE_FROM_PRIMS ? 4 : 2);
    if (width == 4 && bvh.root >= 0) {
      int32_t root4 = B200_EMPTY_ROOT;
      uint32_t depth4 = 0;
      wide = collapse4(bvh, root4, depth4);
      if (3 * depth4 <= B200_STACK) { // up to three pushes per level; deeper trees keep the BVH2
        bvh.root = root4;
        n_nodes4 = (uint32_t)wide.size();
        bvh.max_depth = depth4;
      } else {
        wide.clear();
      }
    }
  }
  // triangle vertices as binary32 rows when every coordinate is exactly representable (glTF positions are: src/main.cc:361-365):
  // 48 bytes per triangle test instead of 80, widened after the load, so the binary64 arithmetic sees identical operands
  std::vector<float> tri32;
  if (c->tri_f32 && d->n_tris) {
    tri32.resize((size_t)d->n_tris * 12, 0.0f);
    bool exact = true;
    for (size_t i = 0; i < d->n_tris && exact; i++)
      for (int k = 0; k < 9; k++) {
        const double v = d->tris[i * B200RT_TRI_STRIDE + k];
        const float f = (float)v;
        if (!((double)f == v)) { // also rejects NaN
          exact = false;
          break;
        }
        tri32[i * 12 + k] = f;
      }
    if (!exact) tri32.clear();
  }
  {
    auto pad = [](size_t n) { return ((n ? n : 16) + 255) & ~(size_t)255; };
    size_t total = 0;
    total += pad(wide.size() * sizeof(BvhNode4)) + pad(bvh.nodes.size() * sizeof(BvhNode)) + pad(ordered.size() * sizeof(Entry)) + pad(ordered.size());
    total += pad(tri32.size() * 4) + pad((size_t)d->n_tris * B200RT_TRI_STRIDE * 8) + pad((size_t)d->n_spheres * B200RT_SPHERE_STRIDE * 8) + pad((size_t)d->n_quads * B200RT_QUAD_STRIDE * 8);
    total += pad(fs.chain_ops.size() * sizeof(ChainOp)) + pad(fs.chains.size() * sizeof(Chain)) + pad(fs.vol_inst.size() * sizeof(VolInst)) + pad(fs.bentries.size() * sizeof(Entry));
    total += pad((size_t)d->n_spheres * 4) + pad((size_t)d->n_quads * 4) + pad((size_t)d->n_tris * 4) + pad(d->n_volumes * sizeof(b200rt_volume));
    total += pad(d->n_materials * sizeof(b200rt_material)) + pad(d->n_textures * sizeof(b200rt_texture));
    CU(c->scene_mem.reserve(total)); // images, noise tables and the reference graph (mirror traversal only) stay outside
  }
  // ---- hot: read per node / per leaf by the traversal
  if (!use_device_builder) {
    if (!wide.empty())
      UP(wide.data(), wide.size(), &S.bvh4, "bvh4");
    else
      UP(bvh.nodes.data(), bvh.nodes.size(), &S.bvh, "bvh");
    UP(ordered.data(), ordered.size(), &S.entries, "entries");
  }
  if (!tri32.empty()) {
    const float *p32 = nullptr;
    UP(tri32.data(), tri32.size(), &p32, "tri32");
    S.tri32 = reinterpret_cast<const float4 *>(p32);
  }
  UP(d->spheres, (size_t)d->n_spheres * B200RT_SPHERE_STRIDE, &S.spheres, "spheres");
  UP(d->quads, (size_t)d->n_quads * B200RT_QUAD_STRIDE, &S.quads, "quads");
  if (tri32.empty()) UP(d->tris, (size_t)d->n_tris * B200RT_TRI_STRIDE, &S.tris, "tris");
  UP(fs.chain_ops.data(), fs.chain_ops.size(), &S.chain_ops, "chain_ops");
  UP(fs.chains.data(), fs.chains.size(), &S.chains, "chains");
  c->hot_bytes = c->scene_mem.arena_used;
  if (c->l2_persist && c->l2_persist_max) { // L2 set-aside for the persisting window (device-wide limit; rounded by the driver)
    cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, std::min(c->l2_persist_max, (c->hot_bytes + ((size_t)1 << 20)) & ~(((size_t)1 << 20) - 1)));
    cudaGetLastError();
  }
  // ---- warm: read per finished ray / per shaded hit
  if (!tri32.empty()) UP(d->tris, (size_t)d->n_tris * B200RT_TRI_STRIDE, &S.tris, "tris"); // normals for finish_hit
  UP(d->sphere_mat, d->n_spheres, &S.sphere_mat, "sphere_mat");
  UP(d->quad_mat, d->n_quads, &S.quad_mat, "quad_mat");
  UP(d->tri_mat, d->n_tris, &S.tri_mat, "tri_mat");
  if (d->tri_uv) UP(d->tri_uv, (size_t)d->n_tris * 6, &S.tri_uv, "tri_uv");
  UP(d->volumes, d->n_volumes, &S.volumes, "volumes");
  UP(d->materials, d->n_materials, &S.materials, "materials");
  UP(d->textures, d->n_textures, &S.textures, "textures");
  UP(fs.vol_inst.data(), fs.vol_inst.size(), &S.vol_inst, "vol_inst");
  UP(fs.bentries.data(), fs.bentries.size(), &S.bentries, "bentries");
  // ---- cold
  UP(d->nodes, d->n_nodes, &S.nodes, "nodes");
  UP(d->list_children, d->n_list_children, &S.list_children, "list_children");
  UP(d->images, d->n_images, &S.images, "images");
  UP(d->image_bytes, d->image_bytes_size, &S.image_bytes, "image_bytes");
  UP(d->perlin_vec, (size_t)d->n_perlin * 768, &S.perlin_vec, "perlin_vec");
  UP(d->perlin_perm, (size_t)d->n_perlin * 256, &S.perlin_perm, "perlin_perm");
  UP(d->value_noise, d->n_value_noise, &S.value_noise, "value_noise");
  UP(d->value_lattice, d->value_lattice_size, &S.value_lattice, "value_lattice");
  S.light_kind = d->light_kind;
  S.light_index = d->light_index;
  S.background_tex = d->background_tex;
  if (use_device_builder) {
    rc = build_bvh_device(c, fs, S, c->bvh_info);
    if (rc != B200RT_OK) return rc;
  }
  if (!use_device_builder) {
    S.bvh_root = bvh.root;
    S.n_bvh_nodes = (uint32_t)bvh.nodes.size();
    S.n_entries = (uint32_t)ordered.size();
    for (int k = 0; k < 3; k++) S.bound[k] = bvh.bound[k];
  }
  S.has_volumes = fs.has_volumes ? 1 : 0;
  { // per entry: the shade queue of its material, derived on the device from what has just been uploaded
    uint8_t *eq = nullptr;
    CU(c->scene_mem.alloc((void **)&eq, S.n_entries));
    if (S.n_entries) k_fill_entry_queue<<<(int)((S.n_entries + 255) / 256), 256, 0, c->stream>>>(S, eq);
    CU(cudaGetLastError());
    S.entry_queue = eq;
  }
  { // the structure itself in global memory, for out-of-line device functions (DevScene::self)
    DevScene *self = nullptr;
    CU(c->scene_mem.alloc((void **)&self, sizeof(DevScene)));
    S.self = self;
    CU(cudaMemcpyAsync(self, &S, sizeof(DevScene), cudaMemcpyHostToDevice, c->stream));
  }
  CU(cudaStreamSynchronize(c->stream)); // host vectors go out of scope
  c->S = S;
  if (!use_device_builder) {
    c->bvh_info.n_nodes = n_nodes4 ? n_nodes4 : (uint32_t)bvh.nodes.size(); // the array that is traversed
    c->bvh_info.n_leaves = bvh.n_leaves;
    c->bvh_info.n_prims = (uint32_t)ordered.size();
    c->bvh_info.max_depth = bvh.max_depth;
    c->bvh_info.build_ms = bvh.build_ms;
    c->bvh_info.sah_cost = bvh.sah_cost;
  }
  c->graph_depth = fs.graph_depth;
  c->n_materials = d->n_materials;
  c->kinds_present = 0;
  for (uint32_t i = 0; i < d->n_materials; i++) c->kinds_present |= 1u << d->materials[i].kind;
  c->bg_mode = d->background_tex < 0 ? 1 : 0;
  { // triangle meshes behind a BVH4 run the warp-uniform extend loop unless the caller says otherwise (B200RT_OPT_EXTEND_LOOP)
    size_t n_tri_entries = 0;
    for (const Entry &e : fs.entries) n_tri_entries += entry_kind(e) == E_TRI;
    c->mesh_scene = S.bvh4 != nullptr && n_tri_entries * 10 >= fs.entries.size() * 9;
  }
  c->scene_optins = d->tri_uv != nullptr;
  for (uint32_t i = 0; i < d->n_textures; i++) c->scene_optins |= d->textures[i].kind == B200RT_TEX_PICTURE_LINEAR;
  for (uint32_t i = 0; i < d->n_images; i++) c->scene_optins |= d->images[i].bytes_per_pixel >= 12;
  c->have_scene = true;
  return B200RT_OK;
}

int b200rt_bvh_info_get(b200rt_ctx *c, b200rt_bvh_info *out) {
  if (!c || !out) return fail(B200RT_ERR_INVALID, "NULL argument");
  if (!c->have_scene) return fail(B200RT_ERR_STATE, "no scene uploaded");
  *out = c->bvh_info;
  return B200RT_OK;
}

int b200rt_set_camera(b200rt_ctx *c, const b200rt_camera *cam) {
  if (!c || !cam) return fail(B200RT_ERR_INVALID, "NULL argument");
  if (cam->mode < B200RT_CAM_PERSPECTIVE || cam->mode > B200RT_CAM_LENS) return fail(B200RT_ERR_INVALID, "unknown camera mode %d", cam->mode);
  if (cam->image_width <= 0 || cam->image_height <= 0) return fail(B200RT_ERR_INVALID, "image size %dx%d", cam->image_width, cam->image_height);
  if ((uint64_t)cam->image_width * cam->image_height >= 0xFFFFFFFFull) return fail(B200RT_ERR_INVALID, "image too large");
  c->cam = *cam;
  c->have_camera = true;
  return B200RT_OK;
}

} // extern "C"

namespace {

int grid_for(const b200rt_ctx *c, size_t n, int block, int per_sm) {
  size_t g = (n + block - 1) / block;
  const size_t cap = (size_t)c->sm_count * per_sm;
  if (g > cap) g = cap;
  if (g < 1) g = 1;
  return (int)g;
}

// the mirror traversal recurses on the device: reserve a stack deep enough for this scene's graph
int prepare_mirror(b200rt_ctx *c) {
  const size_t want = 4096 + (size_t)c->graph_depth * 768;
  size_t have = 0;
  CU(cudaDeviceGetLimit(&have, cudaLimitStackSize));
  if (have < want) CU(cudaDeviceSetLimit(cudaLimitStackSize, want));
  return B200RT_OK;
}

struct ScratchFree {
  std::vector<void *> p;
  ~ScratchFree() {
    for (void *q : p) cudaFree(q);
  }
};

} // namespace

extern "C" {

int b200rt_trace(b200rt_ctx *c, const double *rays7, size_t n, double tmin, double tmax, uint64_t seed, int32_t *prim, double *t, double *uv,
                 uint8_t *front, double *p, double *normal) {
  if (!c) return fail(B200RT_ERR_INVALID, "context is NULL");
  if (!c->have_scene) return fail(B200RT_ERR_STATE, "b200rt_trace before b200rt_upload_scene");
  if (n == 0) return B200RT_OK;
  if (!rays7) return fail(B200RT_ERR_INVALID, "rays7 is NULL");
  CU(cudaSetDevice(c->device));
  ScratchFree sf;
  auto dalloc = [&](void **q, size_t bytes) {
    cudaError_t e = cudaMalloc(q, bytes);
    if (e == cudaSuccess) sf.p.push_back(*q);
    return e;
  };
  double *d_rays = nullptr, *d_t = nullptr, *d_uv = nullptr, *d_p = nullptr, *d_n = nullptr;
  int32_t *d_prim = nullptr;
  uint8_t *d_front = nullptr;
  CU(dalloc((void **)&d_rays, n * 7 * sizeof(double)));
  CU(cudaMemcpyAsync(d_rays, rays7, n * 7 * sizeof(double), cudaMemcpyHostToDevice, c->stream));
  if (prim) CU(dalloc((void **)&d_prim, n * sizeof(int32_t)));
  if (t) CU(dalloc((void **)&d_t, n * sizeof(double)));
  if (uv) CU(dalloc((void **)&d_uv, n * 2 * sizeof(double)));
  if (front) CU(dalloc((void **)&d_front, n));
  if (p) CU(dalloc((void **)&d_p, n * 3 * sizeof(double)));
  if (normal) CU(dalloc((void **)&d_n, n * 3 * sizeof(double)));
  const int grid = grid_for(c, n, 128, 16);
  if (c->traversal == B200RT_TRAVERSAL_MIRROR) {
    int rc = prepare_mirror(c);
    if (rc != B200RT_OK) return rc;
    k_trace<true><<<grid, 128, 0, c->stream>>>(c->S, d_rays, n, tmin, tmax, seed, d_prim, d_t, d_uv, d_front, d_p, d_n);
  } else {
    k_trace<false><<<grid, 128, 0, c->stream>>>(c->S, d_rays, n, tmin, tmax, seed, d_prim, d_t, d_uv, d_front, d_p, d_n);
  }
  CU(cudaGetLastError());
  if (prim) CU(cudaMemcpyAsync(prim, d_prim, n * sizeof(int32_t), cudaMemcpyDeviceToHost, c->stream));
  if (t) CU(cudaMemcpyAsync(t, d_t, n * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
  if (uv) CU(cudaMemcpyAsync(uv, d_uv, n * 2 * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
  if (front) CU(cudaMemcpyAsync(front, d_front, n, cudaMemcpyDeviceToHost, c->stream));
  if (p) CU(cudaMemcpyAsync(p, d_p, n * 3 * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
  if (normal) CU(cudaMemcpyAsync(normal, d_n, n * 3 * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
  CU(cudaStreamSynchronize(c->stream));
  return B200RT_OK;
}

int b200rt_texture_sample(b200rt_ctx *c, int tex, const double *uvp5, size_t n, double *rgb) {
  if (!c) return fail(B200RT_ERR_INVALID, "context is NULL");
  if (!c->have_scene) return fail(B200RT_ERR_STATE, "b200rt_texture_sample before b200rt_upload_scene");
  if (n == 0) return B200RT_OK;
  if (!uvp5 || !rgb) return fail(B200RT_ERR_INVALID, "NULL buffer");
  CU(cudaSetDevice(c->device));
  ScratchFree sf;
  double *d_in = nullptr, *d_out = nullptr;
  CU(cudaMalloc((void **)&d_in, n * 5 * sizeof(double)));
  sf.p.push_back(d_in);
  CU(cudaMalloc((void **)&d_out, n * 3 * sizeof(double)));
  sf.p.push_back(d_out);
  CU(cudaMemcpyAsync(d_in, uvp5, n * 5 * sizeof(double), cudaMemcpyHostToDevice, c->stream));
  k_texture<<<grid_for(c, n, 128, 16), 128, 0, c->stream>>>(c->S, tex, d_in, n, d_out);
  CU(cudaGetLastError());
  CU(cudaMemcpyAsync(rgb, d_out, n * 3 * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
  CU(cudaStreamSynchronize(c->stream));
  return B200RT_OK;
}

int b200rt_scatter(b200rt_ctx *c, const double *rays7, const double *rec12, const double *draws6, size_t n, double *out16) {
  if (!c) return fail(B200RT_ERR_INVALID, "context is NULL");
  if (!c->have_scene) return fail(B200RT_ERR_STATE, "b200rt_scatter before b200rt_upload_scene");
  if (n == 0) return B200RT_OK;
  if (!rays7 || !rec12 || !draws6 || !out16) return fail(B200RT_ERR_INVALID, "NULL buffer");
  for (size_t i = 0; i < n; i++) {
    const double m = rec12[12 * i + 10];
    if (!(m >= 0 && m < (double)c->n_materials) || m != (double)(int)m) return fail(B200RT_ERR_INVALID, "b200rt_scatter: record %zu names material %g (scene has %u)", i, m, c->n_materials);
  }
  CU(cudaSetDevice(c->device));
  DevMem tmp;
  double *d_r = nullptr, *d_h = nullptr, *d_u = nullptr, *d_o = nullptr;
  CU(tmp.alloc((void **)&d_r, n * 7 * sizeof(double)));
  CU(tmp.alloc((void **)&d_h, n * 12 * sizeof(double)));
  CU(tmp.alloc((void **)&d_u, n * 6 * sizeof(double)));
  CU(tmp.alloc((void **)&d_o, n * 16 * sizeof(double)));
  CU(cudaMemcpyAsync(d_r, rays7, n * 7 * sizeof(double), cudaMemcpyHostToDevice, c->stream));
  CU(cudaMemcpyAsync(d_h, rec12, n * 12 * sizeof(double), cudaMemcpyHostToDevice, c->stream));
  CU(cudaMemcpyAsync(d_u, draws6, n * 6 * sizeof(double), cudaMemcpyHostToDevice, c->stream));
  k_scatter<<<grid_for(c, n, 128, 16), 128, 0, c->stream>>>(c->S, d_r, d_h, d_u, n, d_o);
  CU(cudaGetLastError());
  CU(cudaMemcpyAsync(out16, d_o, n * 16 * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
  CU(cudaStreamSynchronize(c->stream));
  return B200RT_OK;
}

int b200rt_generate_rays(b200rt_ctx *c, uint64_t seed, int spp_begin, int spp_end, double *rays7) {
  if (!c || !rays7) return fail(B200RT_ERR_INVALID, "NULL argument");
  if (!c->have_camera) return fail(B200RT_ERR_STATE, "b200rt_generate_rays before b200rt_set_camera");
  if (spp_end < spp_begin || spp_begin < 0) return fail(B200RT_ERR_INVALID, "bad sample range [%d, %d)", spp_begin, spp_end);
  const size_t n = (size_t)c->cam.image_width * c->cam.image_height * (size_t)(spp_end - spp_begin);
  if (n == 0) return B200RT_OK;
  CU(cudaSetDevice(c->device));
  ScratchFree sf;
  double *d = nullptr;
  CU(cudaMalloc((void **)&d, n * 7 * sizeof(double)));
  sf.p.push_back(d);
  k_raygen<<<grid_for(c, n, 256, 8), 256, 0, c->stream>>>(make_camera(c->cam), seed, spp_begin, n, d);
  CU(cudaGetLastError());
  CU(cudaMemcpyAsync(rays7, d, n * 7 * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
  CU(cudaStreamSynchronize(c->stream));
  return B200RT_OK;
}

} // extern "C"

namespace {

int ensure_wave(b200rt_ctx::WaveSet &w, uint32_t capacity, bool mirror, bool sort = false) {
  if (w.capacity >= capacity && (!mirror || w.mirror) && (!sort || w.sort_tmp)) return B200RT_OK;
  w.drop_graph(); // its kernel nodes point into the pools that are about to be freed
  w.mem.release();
  w.capacity = 0;
  auto col = [&](void **p, size_t elem) { return w.mem.alloc(p, (size_t)capacity * elem); };
  for (int b = 0; b < 2; b++) {
    CU(col((void **)&w.pools.b[b].base, B200_PATH_QUADS * sizeof(double2)));
    w.pools.b[b].capacity = capacity;
  }
  CU(col((void **)&w.hits, sizeof(HitRef)));
  w.recs = nullptr;
  if (mirror) CU(col((void **)&w.recs, sizeof(Rec)));
  for (int k = 0; k < B200_NQUEUES; k++) CU(col((void **)&w.queues.q[k], 4));
  CU(w.mem.alloc((void **)&w.ctl, sizeof(WaveCtl)));
  w.sort_tmp = nullptr;
  if (sort) {
    for (int k = 0; k < 2; k++) {
      CU(col((void **)&w.sort_keys[k], 4));
      CU(col((void **)&w.sort_vals[k], 4));
    }
    w.sort_tmp_bytes = 0;
    CU(cub::DeviceRadixSort::SortPairs(nullptr, w.sort_tmp_bytes, w.sort_keys[0], w.sort_keys[1], w.sort_vals[0], w.sort_vals[1], (int)capacity, 0, 21));
    CU(w.mem.alloc(&w.sort_tmp, w.sort_tmp_bytes));
  }
  w.capacity = capacity;
  w.mirror = mirror;
  return B200RT_OK;
}

// Persistent extend launch.  The shared-memory carve-out is requested explicitly (smallest configuration that holds the
// resident CTAs): what is left of the 256 KB is the L1 that caches BVH nodes and primitives.
template <class K>
void launch_extend(K kernel, b200rt_ctx *c, b200rt_ctx::WaveSet &w, int ctas_per_sm, int threads, int smem_per_cta, cudaStream_t s, const PathBuf &cur, const WaveParams &wp) {
  const int want_kb = (ctas_per_sm * (smem_per_cta + 1024) + 1023) / 1024;
  cudaFuncSetAttribute(kernel, cudaFuncAttributePreferredSharedMemoryCarveout, c->carveout_pct >= 0 ? c->carveout_pct : std::min(100, (want_kb * 100 + 227) / 228));
  kernel<<<c->sm_count * ctas_per_sm, threads, 0, s>>>(w.ctl, c->S, cur, w.hits, wp, (const uint32_t *)(c->ray_sort ? w.sort_vals[1] : nullptr));
}

template <bool MIRROR>
int launch_shades(b200rt_ctx *c, b200rt_ctx::WaveSet &w, cudaStream_t s, int grid, const PathBuf &cur, const PathBuf &nxt, double *accum, const WaveParams &wp) {
  int n = 0;
#define SHADE(K)                                                                                                       \
  if (K == B200_Q_MISS ? wp.fused_miss == 0 : (c->kinds_present >> K & 1u))                                                                \
  n++, (c->scene_optins ? k_shade<K, MIRROR, true> : k_shade<K, MIRROR, false>)<<<grid, 128, 0, s>>>(w.ctl, c->S, cur, nxt, w.hits, w.recs, w.queues.q[K], accum, wp)
  SHADE(B200_Q_MISS);
  SHADE(B200RT_MAT_DIFFUSE_LIGHT);
  SHADE(B200RT_MAT_LAMBERTIAN);
  SHADE(B200RT_MAT_METAL);
  SHADE(B200RT_MAT_DIELECTRIC);
  SHADE(B200RT_MAT_GLOSS);
  SHADE(B200RT_MAT_ISOTROPIC);
#undef SHADE
  return n;
}

// One wave of one pipeline on stream s: generate -> extend -> classify -> shade per queue -> roll the counters over.
// Returns the number of launches (negative: error).  ext_ctas = resident extend CTAs per SM (4 alone; fewer when a second
// pipeline's shade kernels are meant to run beside it).  ext_events (may be NULL) receives an event pair around extend.
// `more` (may be 0): the conditional handle of the graph while loop this wave is being captured into; the last kernel then
// also decides whether the loop goes round again.
int launch_wave(b200rt_ctx *c, b200rt_ctx::WaveSet &w, cudaStream_t s, unsigned long long wave, const DevCamera &cam, const WaveParams &wp, int g_gen, int g_ext,
                double *d_rgb_sum, int ext_ctas, std::vector<cudaEvent_t> *ext_events, cudaGraphConditionalHandle more = 0) {
  const bool mirror = c->traversal == B200RT_TRAVERSAL_MIRROR;
  const PathBuf &cur = w.pools.b[wave & 1], &nxt = w.pools.b[(wave & 1) ^ 1];
  const bool compact = cam.c.mode == B200RT_CAM_FISHEYE; // the only camera with samples that produce no ray
  int n_gen_launches = 2;
  if (compact) {
    // dropped samples leave room in the wave: further passes top it up (each sees the n_cur / next_sample the last one left)
    const int passes = 2; // 1: -2.7 % on Fox, 3: same as 2 (profiles/r01_x_fox_gen_passes.txt)
    for (int k = 0; k < passes; k++) {
#if B200_GENERATE_POOLS
      k_generate<true><<<g_gen, 256, 0, s>>>(w.ctl, w.pools, (int)(wave & 1), cam, wp);
#else
      k_generate<true><<<g_gen, 256, 0, s>>>(w.ctl, cur, cam, wp);
#endif
      k_ctl_generated<<<1, 1, 0, s>>>(w.ctl, 1);
    }
    n_gen_launches = 2 * passes;
  } else {
#if B200_GENERATE_POOLS
    k_generate<false><<<g_gen, 256, 0, s>>>(w.ctl, w.pools, (int)(wave & 1), cam, wp);
#else
    k_generate<false><<<g_gen, 256, 0, s>>>(w.ctl, cur, cam, wp);
#endif
    k_ctl_generated<<<1, 1, 0, s>>>(w.ctl, 0);
  }
  const bool tail = !mirror && c->tail_paths > 0;
  if (tail) (c->scene_optins ? k_tail<true> : k_tail<false>)<<<(int)((c->tail_paths + 63) / 64), 64, 0, s>>>(w.ctl, c->S, cur, d_rgb_sum, wp, c->tail_paths);
  if (c->ray_sort && !mirror) { // experiment: a full radix sort per wave (its cost is not the point: the extend time after it is)
    k_sort_keys<<<g_gen, 256, 0, s>>>(w.ctl, cur, wp, c->ray_sort, w.sort_keys[0], w.sort_vals[0], w.capacity);
    CU(cub::DeviceRadixSort::SortPairs(w.sort_tmp, w.sort_tmp_bytes, w.sort_keys[0], w.sort_keys[1], w.sort_vals[0], w.sort_vals[1], (int)w.capacity, 0, 21, s));
  }
  if (ext_events) {
    cudaEvent_t a, b; // owned by the caller's guard from the moment they exist
    CU(cudaEventCreate(&a));
    ext_events->push_back(a);
    CU(cudaEventCreate(&b));
    ext_events->push_back(b);
    CU(cudaEventRecord(a, s));
  }
  if (mirror)
    k_extend<true><<<g_ext, 128, 0, s>>>(w.ctl, c->S, cur, w.hits, w.recs, w.queues, wp);
  else
#ifndef B200_BVH4_OCCUPANCY_EXPERIMENT
#define B200_BVH4_OCCUPANCY_EXPERIMENT 0 // 1: B200RT_OPT_EXTEND_BLOCKS 5 / 6 also select 96- / 80-register BVH4 kernels (measured: slower)
#endif
    switch (c->S.bvh4 ? (B200_BVH4_OCCUPANCY_EXPERIMENT && c->extend_blocks >= 5 ? (c->loop_mode() == 1 ? 200 : 100) + c->extend_blocks : 0)
                      : c->extend_blocks * 2 + (c->speculative ? 1 : 0)) { // the other tuning variants walk the BVH2 only
    // EXTEND_BLOCKS names the resident WARPS per SM / 4; registers per thread follow (65536 / 32 / warps, multiple of 8)
#define EXT(MAXREG, SPEC, WARPS, STAGES)                                                                               \
  launch_extend(k_extend_persistent<MAXREG, SPEC, WARPS, STAGES, false>, c, w, 65536 / (MAXREG * WARPS * 32), WARPS * 32, WARPS * STAGES * 3584, s, cur, wp)
    case 12: EXT(80, false, 4, 1); break;  // 24 warps per SM
    case 10: EXT(96, false, 2, 1); break;  // 20 warps
    case 18: EXT(112, false, 2, 1); break; // 18 warps (option value 9: 2-warp CTAs)
    case 34: EXT(120, false, 1, 1); break; // 17 warps (option value 17: 1-warp CTAs)
    case 9: EXT(128, true, 4, 2); break;
#undef EXT
#if B200_BVH4_OCCUPANCY_EXPERIMENT
    // BVH4 at higher occupancy (latency-bound scenes: the Sponza stand-in waits on L2 with four warps per scheduler)
    case 105: launch_extend(k_extend_persistent<96, false, 4, 2, true>, c, w, 5, 128, 4 * 2 * 3584, s, cur, wp); break; // 20 warps per SM
    case 106: launch_extend(k_extend_persistent<80, false, 4, 2, true>, c, w, 6, 128, 4 * 2 * 3584, s, cur, wp); break; // 24 warps per SM
    // the same under the warp-uniform loop, and with one staging buffer per warp (more of the 256 KB left as L1)
    case 205: launch_extend(k_extend_persistent<96, false, 4, 2, true, true>, c, w, 5, 128, 4 * 2 * 3584, s, cur, wp); break;
    case 206: launch_extend(k_extend_persistent<80, false, 4, 2, true, true>, c, w, 6, 128, 4 * 2 * 3584, s, cur, wp); break;
    case 214: launch_extend(k_extend_persistent<128, false, 4, 1, true, true>, c, w, ext_ctas, 128, 4 * 1 * 3584, s, cur, wp); break;
    case 215: launch_extend(k_extend_persistent<96, false, 4, 1, true, true>, c, w, 5, 128, 4 * 1 * 3584, s, cur, wp); break;
#endif
    default: // 128 registers, 4-warp CTAs
      if (c->loop_mode() == 1 && c->S.bvh4)
        launch_extend(k_extend_persistent<128, false, 4, 2, true, true>, c, w, ext_ctas, 128, 4 * 2 * 3584, s, cur, wp);
      else if (c->loop_mode() == 1)
        launch_extend(k_extend_persistent<128, false, 4, 2, false, true>, c, w, ext_ctas, 128, 4 * 2 * 3584, s, cur, wp);
      else if (c->S.bvh4)
        launch_extend(k_extend_persistent<128, false, 4, 2, true>, c, w, ext_ctas, 128, 4 * 2 * 3584, s, cur, wp);
      else if (c->extend_stages == 1 && wp.refill == 32)
        launch_extend(k_extend_persistent<128, false, 4, 1, false>, c, w, ext_ctas, 128, 4 * 1 * 3584, s, cur, wp);
      else
        launch_extend(k_extend_persistent<128, false, 4, 2, false>, c, w, ext_ctas, 128, 4 * 2 * 3584, s, cur, wp);
      break;
    }
  if (ext_events) CU(cudaEventRecord(ext_events->back(), s));
  if (!mirror) k_classify<<<g_gen, 256, 0, s>>>(w.ctl, w.hits, w.queues);
  const int n_shade = mirror ? launch_shades<true>(c, w, s, g_ext, cur, nxt, d_rgb_sum, wp) : launch_shades<false>(c, w, s, g_ext, cur, nxt, d_rgb_sum, wp);
  if (more)
    k_ctl_swap_cond<<<1, 1, 0, s>>>(w.ctl, more);
  else
    k_ctl_swap<<<1, 1, 0, s>>>(w.ctl);
  return (mirror ? 2 : 3) + (tail ? 1 : 0) + n_gen_launches + n_shade;
}

// The hot prefix of the scene arena (tree, leaf entries, primitive rows) as a persisting L2 window: what every ray re-reads
// stays in L2 while the wave's path state streams through it.  Returns false when the device / scene has nothing to pin.
bool l2_window(const b200rt_ctx *c, cudaAccessPolicyWindow &w) {
  if (!c->l2_persist || !c->hot_bytes || !c->l2_window_max || !c->l2_persist_max || !c->scene_mem.arena) return false;
  const size_t win = std::min(c->hot_bytes, c->l2_window_max);
  w.base_ptr = c->scene_mem.arena;
  w.num_bytes = win;
  w.hitRatio = (float)std::min(1.0, (double)c->l2_persist_max / (double)win);
  w.hitProp = cudaAccessPropertyPersisting;
  w.missProp = cudaAccessPropertyStreaming;
  return true;
}
void l2_window_on_stream(const b200rt_ctx *c, cudaStream_t s) {
  cudaStreamAttrValue v;
  std::memset(&v, 0, sizeof v);
  if (!l2_window(c, v.accessPolicyWindow)) v.accessPolicyWindow.num_bytes = 0; // a zero-size window switches the policy off
  cudaStreamSetAttribute(s, cudaStreamAttributeAccessPolicyWindow, &v);
  cudaGetLastError(); // a refused hint is not an error of the render
}

// The device-driven wave loop of one pipeline as a CUDA graph: a single conditional WHILE node whose body is one wave
// (captured from launch_wave) and whose last kernel, k_ctl_swap_cond, sets the loop condition from the device-side counters.
// Everything a kernel node bakes in goes into `key`; an unchanged key re-launches the instantiated graph as it is.
int ensure_wave_graph(b200rt_ctx *c, b200rt_ctx::WaveSet &w, const std::string &key, const DevCamera &cam, const WaveParams &wp, int g_gen, int g_ext,
                      double *d_rgb_sum, int ext_ctas) {
  if (w.graph_exec && w.graph_key == key) return B200RT_OK;
  w.drop_graph();
  CU(cudaGraphCreate(&w.graph, 0));
  cudaGraphConditionalHandle more;
  CU(cudaGraphConditionalHandleCreate(&more, w.graph, 1, cudaGraphCondAssignDefault)); // every launch starts with "go round"
  cudaGraphNodeParams np = {cudaGraphNodeTypeConditional};
  np.type = cudaGraphNodeTypeConditional;
  np.conditional.handle = more;
  np.conditional.type = cudaGraphCondTypeWhile;
  np.conditional.size = 1;
  cudaGraphNode_t node;
  CU(cudaGraphAddNode(&node, w.graph, nullptr, 0, &np));
  cudaGraph_t body = np.conditional.phGraph_out[0];
  l2_window_on_stream(c, c->cap_stream);
  CU(cudaStreamBeginCaptureToGraph(c->cap_stream, body, nullptr, nullptr, 0, cudaStreamCaptureModeRelaxed));
  // the body = an even and an odd wave (pools are by-value kernel parameters, so a wave's launches differ with its parity); the
  // second one decides whether the loop goes round again.  A frame that ends on an even wave pays for one empty odd wave.
  int n = launch_wave(c, w, c->cap_stream, 0, cam, wp, g_gen, g_ext, d_rgb_sum, ext_ctas, nullptr, 0);
  if (n >= 0) {
    const int n2 = launch_wave(c, w, c->cap_stream, 1, cam, wp, g_gen, g_ext, d_rgb_sum, ext_ctas, nullptr, more);
    n = n2 < 0 ? n2 : n; // launches per WAVE
  }
  cudaGraph_t captured = nullptr;
  const cudaError_t e_end = cudaStreamEndCapture(c->cap_stream, &captured);
  if (n < 0) {
    w.drop_graph();
    return n;
  }
  if (e_end != cudaSuccess) {
    w.drop_graph();
    return fail(B200RT_ERR_CUDA, "capturing the wave graph: %s", cudaGetErrorString(e_end));
  }
  { // the access-policy window on every kernel node of the body (explicitly: not every driver carries the capturing stream's over)
    cudaLaunchAttributeValue av;
    std::memset(&av, 0, sizeof av);
    if (l2_window(c, av.accessPolicyWindow)) {
      size_t n_nodes = 0;
      if (cudaGraphGetNodes(body, nullptr, &n_nodes) == cudaSuccess && n_nodes) {
        std::vector<cudaGraphNode_t> nodes(n_nodes);
        cudaGraphGetNodes(body, nodes.data(), &n_nodes);
        for (cudaGraphNode_t nd : nodes) {
          cudaGraphNodeType ty;
          if (cudaGraphNodeGetType(nd, &ty) == cudaSuccess && ty == cudaGraphNodeTypeKernel) cudaGraphKernelNodeSetAttribute(nd, cudaLaunchAttributeAccessPolicyWindow, &av);
        }
      }
      cudaGetLastError();
    }
  }
  const cudaError_t e_inst = cudaGraphInstantiate(&w.graph_exec, w.graph, 0);
  if (e_inst != cudaSuccess) {
    w.drop_graph();
    return fail(B200RT_ERR_CUDA, "instantiating the wave graph: %s", cudaGetErrorString(e_inst));
  }
  w.graph_key = key;
  w.graph_launches_per_wave = n;
  return B200RT_OK;
}

template <class T> void key_add(std::string &k, const T &v) { k.append(reinterpret_cast<const char *>(&v), sizeof v); }

} // namespace

extern "C" {

int b200rt_render_device(b200rt_ctx *c, uint64_t seed, int spp_begin, int spp_end, double *d_rgb_sum, void *cuda_stream, b200rt_stats *stats) {
  if (!c || !d_rgb_sum) return fail(B200RT_ERR_INVALID, "NULL argument");
  if (!c->have_scene) return fail(B200RT_ERR_STATE, "render before b200rt_upload_scene");
  if (!c->have_camera) return fail(B200RT_ERR_STATE, "render before b200rt_set_camera");
  if (spp_end < spp_begin || spp_begin < 0) return fail(B200RT_ERR_INVALID, "bad sample range [%d, %d)", spp_begin, spp_end);
  CU(cudaSetDevice(c->device));
  cudaStream_t s = (cudaStream_t)cuda_stream;
  const uint32_t npix = (uint32_t)c->cam.image_width * (uint32_t)c->cam.image_height;
  const unsigned long long total = (unsigned long long)npix * (unsigned long long)(spp_end - spp_begin);
  if (stats) std::memset(stats, 0, sizeof *stats);
  if (total == 0) return B200RT_OK;
  if (c->cam.max_depth <= 0) { // ray_color(iteration <= 0) returns color(0) before it traces anything (src/camera.h:194-195): every sample adds 0
    if (stats) stats->samples = total;
    return B200RT_OK;
  }
  const bool mirror = c->traversal == B200RT_TRAVERSAL_MIRROR;
  // Two pipelines: the sample range is split in two halves that run as independent wavefronts on two streams, so that one
  // half's kernels fill the SMs that the other half's kernel has already drained (every launch ends with a tail of
  // partly idle SMs; the persistent extend kernel most of all).  pipe_extend_ctas < 4 would leave registers for the other
  // half's shade kernels to run BESIDE extend; measured slower (extend needs its occupancy), so the default is 4.
  int n_pipe = (c->pipelines > 1 && !mirror && total >= (1ull << 21)) ? std::min(c->pipelines, spp_end - spp_begin) : 1;
  int rc;
  if (mirror && (rc = prepare_mirror(c)) != B200RT_OK) return rc;

  WaveParams wp[B200_MAX_PIPELINES];
  std::memset(wp, 0, sizeof wp); // padding bytes too: the structure is part of the wave graph's cache key
  unsigned long long tot[B200_MAX_PIPELINES] = {0};
  uint32_t cap[B200_MAX_PIPELINES] = {0};
  for (int p = 0; p < n_pipe; p++) {
    // contiguous, disjoint, covering sample sub-ranges whose sizes differ by at most one
    const int span = spp_end - spp_begin, base = span / n_pipe, extra = span % n_pipe;
    const int lo = spp_begin + p * base + std::min(p, extra), hi = lo + base + (p < extra ? 1 : 0);
    WaveParams &q = wp[p];
    q.seed = seed;
    q.npix = npix;
    q.width = (uint32_t)c->cam.image_width;
    q.spp_begin = lo;
    q.max_depth = c->cam.max_depth;
    q.kill_zero = c->kill_zero;
    q.fused_miss = (mirror || !c->fuse_miss) ? 0 : c->bg_mode;
    q.leaf_lanes = c->loop_lanes();
    q.octant_sort = c->octant_sort;
    q.refill = c->refill_lanes ? c->refill_lanes : (c->S.n_bvh_nodes >= 32768 ? 8 : 32);
    for (int k = 0; k < 3; k++) q.bound[k] = c->S.bound[k];
    tot[p] = (unsigned long long)npix * (unsigned long long)(hi - lo);
    uint32_t capacity = c->wave_paths / (uint32_t)n_pipe;
    if (tot[p] < capacity) capacity = (uint32_t)((tot[p] + 31) & ~31ull);
    if ((rc = ensure_wave(c->w[p], capacity, mirror, c->ray_sort != 0 && !mirror)) != B200RT_OK) return rc;
    capacity = c->w[p].capacity; // a larger pool from an earlier call is kept; use min(wave_paths share, pool)
    if (capacity > c->wave_paths / (uint32_t)n_pipe) capacity = c->wave_paths / (uint32_t)n_pipe;
    cap[p] = capacity;
  }
  const DevCamera cam = make_camera(c->cam);
  const int ext_ctas = n_pipe > 1 ? c->pipe_extend_ctas : 4;
  cudaStream_t ps[B200_MAX_PIPELINES] = {s, s, s, s};
  if (n_pipe > 1) {
    CU(cudaEventRecord(c->ev_fork, s));
    for (int p = 0; p < n_pipe; p++) {
      ps[p] = c->w[p].stream;
      CU(cudaStreamWaitEvent(ps[p], c->ev_fork, 0));
    }
  }

  const bool use_graph = c->wave_loop == 1 && !mirror;
  for (int p = 0; p < n_pipe; p++) l2_window_on_stream(c, ps[p]);
  // Whatever happens below (a failed launch, a failed copy), the forked pipeline streams are joined back into the caller's
  // stream and the per-launch timing events are destroyed: the caller may keep using `s` and d_rgb_sum after an error.
  struct WaveLoopGuard {
    b200rt_ctx *c;
    cudaStream_t s;
    cudaStream_t *ps;
    int n_pipe;
    std::vector<cudaEvent_t> ext_events;
    bool joined = false;
    cudaError_t join() {
      if (joined) return cudaSuccess;
      joined = true;
      cudaError_t first = cudaSuccess;
      if (n_pipe > 1)
        for (int p = 0; p < n_pipe; p++) {
          cudaError_t e = cudaEventRecord(c->w[p].done, ps[p]);
          if (e == cudaSuccess) e = cudaStreamWaitEvent(s, c->w[p].done, 0);
          if (e != cudaSuccess && first == cudaSuccess) first = e;
        }
      return first;
    }
    ~WaveLoopGuard() {
      join();
      for (cudaEvent_t e : ext_events) cudaEventDestroy(e);
    }
  } guard{c, s, ps, n_pipe};
  std::vector<cudaEvent_t> &ext_events = guard.ext_events;
  if (stats) CU(cudaEventRecord(c->ev0, s));
  unsigned long long launches = 0, wave = 0, next_check = 0;
  bool done[B200_MAX_PIPELINES];
  for (int p = 0; p < B200_MAX_PIPELINES; p++) done[p] = p >= n_pipe;
  for (int p = 0; p < n_pipe; p++) {
    k_ctl_init<<<1, 1, 0, ps[p]>>>(c->w[p].ctl, tot[p], cap[p]);
    launches++;
    next_check = std::max(next_check, (tot[p] + cap[p] - 1) / cap[p]);
  }
  if (use_graph) {
    // device-driven: one graph launch per pipeline; the while node repeats the wave until the counters say the frame is done
    for (int p = 0; p < n_pipe; p++) {
      std::string key;
      key_add(key, c->S), key_add(key, cam), key_add(key, wp[p]), key_add(key, cap[p]), key_add(key, d_rgb_sum), key_add(key, ext_ctas);
      key_add(key, c->tail_paths), key_add(key, c->extend_blocks), key_add(key, c->speculative), key_add(key, c->extend_stages), key_add(key, c->carveout_pct), key_add(key, c->loop_mode()), key_add(key, c->l2_persist), key_add(key, c->ray_sort);
      key_add(key, c->kinds_present), key_add(key, c->scene_optins), key_add(key, c->w[p].pools), key_add(key, c->w[p].hits);
      if ((rc = ensure_wave_graph(c, c->w[p], key, cam, wp[p], grid_for(c, cap[p], 256, 8), grid_for(c, cap[p], 128, 16), d_rgb_sum, ext_ctas)) != B200RT_OK) return rc;
      CU(cudaGraphLaunch(c->w[p].graph_exec, ps[p]));
      if (stats) CU(cudaMemcpyAsync(c->w[p].h_ctl, c->w[p].ctl, sizeof(WaveCtl), cudaMemcpyDeviceToHost, ps[p]));
    }
    if (stats)
      for (int p = 0; p < n_pipe; p++) {
        CU(cudaStreamSynchronize(ps[p]));
        launches += c->w[p].h_ctl->waves * (unsigned long long)c->w[p].graph_launches_per_wave;
        wave = std::max(wave, c->w[p].h_ctl->waves);
      }
  } else
  for (;;) {
    for (int p = 0; p < n_pipe; p++) {
      if (done[p]) continue;
      const int n = launch_wave(c, c->w[p], ps[p], wave, cam, wp[p], grid_for(c, cap[p], 256, 8), grid_for(c, cap[p], 128, 16), d_rgb_sum, ext_ctas,
                                stats && n_pipe == 1 ? &ext_events : nullptr);
      if (n < 0) return n;
      launches += (unsigned long long)n;
    }
    CU(cudaGetLastError()); // launch-configuration errors surface here, once per wave
    wave++;
    if (wave >= next_check) {
      for (int p = 0; p < n_pipe; p++)
        if (!done[p]) CU(cudaMemcpyAsync(c->w[p].h_ctl, c->w[p].ctl, sizeof(WaveCtl), cudaMemcpyDeviceToHost, ps[p]));
      unsigned long long ahead = 2;
      for (int p = 0; p < n_pipe; p++) {
        if (done[p]) continue;
        CU(cudaStreamSynchronize(ps[p]));
        const WaveCtl &h = *c->w[p].h_ctl;
        if (h.n_cur == 0 && h.next_sample >= h.total_samples) {
          done[p] = true;
          continue;
        }
        const unsigned long long left = h.total_samples - h.next_sample;
        if (left > 0) ahead = std::max(ahead, std::max<unsigned long long>(1, left / cap[p]));
      }
      if (done[0] && done[1] && done[2] && done[3]) break;
      next_check = wave + ahead;
    }
  }
  CU(guard.join()); // the caller's stream continues when every part is in the accumulator
  if (stats) {
    CU(cudaEventRecord(c->ev1, s));
    CU(cudaEventSynchronize(c->ev1));
    float ms = 0;
    CU(cudaEventElapsedTime(&ms, c->ev0, c->ev1));
    stats->gpu_ms = ms;
    double ext = 0;
    for (size_t k = 0; k + 1 < ext_events.size(); k += 2) {
      float e = 0;
      cudaEventElapsedTime(&e, ext_events[k], ext_events[k + 1]);
      ext += e;
    }
    stats->extend_ms = ext; // not measured with two pipelines (the kernels of the halves overlap)
    stats->samples = total;
    for (int p = 0; p < n_pipe; p++) stats->rays += c->w[p].h_ctl->rays, stats->nan_samples += c->w[p].h_ctl->nan_samples;
    stats->kernel_launches = launches;
    stats->waves = wave;
  }
  return B200RT_OK;
}

int b200rt_render(b200rt_ctx *c, uint64_t seed, int spp_begin, int spp_end, double *rgb_sum, b200rt_stats *stats) {
  if (!c || !rgb_sum) return fail(B200RT_ERR_INVALID, "NULL argument");
  if (!c->have_camera) return fail(B200RT_ERR_STATE, "render before b200rt_set_camera");
  CU(cudaSetDevice(c->device));
  const size_t elems = (size_t)c->cam.image_width * c->cam.image_height * 3;
  if (c->accum_elems < elems) {
    if (c->accum) cudaFree(c->accum);
    c->accum = nullptr, c->accum_elems = 0;
    CU(cudaMalloc((void **)&c->accum, elems * sizeof(double)));
    c->accum_elems = elems;
  }
  CU(cudaMemsetAsync(c->accum, 0, elems * sizeof(double), c->stream));
  int rc = b200rt_render_device(c, seed, spp_begin, spp_end, c->accum, c->stream, stats);
  if (rc != B200RT_OK) return rc;
  CU(cudaMemcpyAsync(rgb_sum, c->accum, elems * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
  CU(cudaStreamSynchronize(c->stream));
  if (stats) {
    stats->h2d_bytes = c->scene_h2d_bytes + sizeof(b200rt_camera);
    stats->d2h_bytes = (double)(elems * sizeof(double));
  }
  return B200RT_OK;
}

int b200rt_finalize(b200rt_ctx *c, const double *rgb_sum, int total_spp, double *rgb_linear, int32_t *ppm_rgb) {
  if (!c || !rgb_sum) return fail(B200RT_ERR_INVALID, "NULL argument");
  if (!c->have_camera) return fail(B200RT_ERR_STATE, "finalize before b200rt_set_camera");
  if (total_spp <= 0) return fail(B200RT_ERR_INVALID, "total_spp must be positive");
  CU(cudaSetDevice(c->device));
  const size_t n = (size_t)c->cam.image_width * c->cam.image_height * 3;
  ScratchFree sf;
  double *d_sum = nullptr, *d_lin = nullptr;
  int32_t *d_ppm = nullptr;
  CU(cudaMalloc((void **)&d_sum, n * sizeof(double)));
  sf.p.push_back(d_sum);
  CU(cudaMemcpyAsync(d_sum, rgb_sum, n * sizeof(double), cudaMemcpyHostToDevice, c->stream));
  if (rgb_linear) {
    CU(cudaMalloc((void **)&d_lin, n * sizeof(double)));
    sf.p.push_back(d_lin);
  }
  if (ppm_rgb) {
    CU(cudaMalloc((void **)&d_ppm, n * sizeof(int32_t)));
    sf.p.push_back(d_ppm);
  }
  k_finalize<<<grid_for(c, n, 256, 8), 256, 0, c->stream>>>(d_sum, n, (double)total_spp, d_lin, d_ppm);
  CU(cudaGetLastError());
  if (rgb_linear) CU(cudaMemcpyAsync(rgb_linear, d_lin, n * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
  if (ppm_rgb) CU(cudaMemcpyAsync(ppm_rgb, d_ppm, n * sizeof(int32_t), cudaMemcpyDeviceToHost, c->stream));
  CU(cudaStreamSynchronize(c->stream));
  return B200RT_OK;
}

size_t b200rt_ppm_bound(int image_width, int image_height) {
  if (image_width <= 0 || image_height <= 0) return 0;
  return 64 + (size_t)image_width * (size_t)image_height * 36; // header + three 11-character integers, two blanks and a newline per pixel
}

int b200rt_format_ppm(b200rt_ctx *c, const double *rgb_sum, int total_spp, char *out, size_t capacity, size_t *out_len) {
  if (!c || !rgb_sum || !out || !out_len) return fail(B200RT_ERR_INVALID, "NULL argument");
  if (!c->have_camera) return fail(B200RT_ERR_STATE, "format_ppm before b200rt_set_camera");
  if (total_spp <= 0) return fail(B200RT_ERR_INVALID, "total_spp must be positive");
  CU(cudaSetDevice(c->device));
  const size_t npix = (size_t)c->cam.image_width * c->cam.image_height;
  char header[64];
  const int hl = snprintf(header, sizeof header, "P3\n%d %d\n255\n", c->cam.image_width, c->cam.image_height); // src/camera.h:149-151
  const size_t bound = b200rt_ppm_bound(c->cam.image_width, c->cam.image_height);
  const size_t dev_cap = capacity < bound ? capacity : bound;
  if (dev_cap < (size_t)hl) return fail(B200RT_ERR_INVALID, "format_ppm: capacity %zu is smaller than the header", capacity);
  const uint32_t n_cta = (uint32_t)((npix + B200_PPM_CTA * B200_PPM_PER_THREAD - 1) / (B200_PPM_CTA * B200_PPM_PER_THREAD));
  DevMem tmp;
  double *d_sum = nullptr;
  int32_t *d_rgb = nullptr;
  unsigned long long *d_cta = nullptr, *d_total = nullptr;
  char *d_out = nullptr;
  CU(tmp.alloc((void **)&d_sum, npix * 3 * sizeof(double)));
  CU(tmp.alloc((void **)&d_rgb, npix * 3 * sizeof(int32_t)));
  CU(tmp.alloc((void **)&d_cta, (size_t)n_cta * 8));
  CU(tmp.alloc((void **)&d_total, 8));
  CU(tmp.alloc((void **)&d_out, dev_cap));
  CU(cudaMemcpyAsync(d_sum, rgb_sum, npix * 3 * sizeof(double), cudaMemcpyHostToDevice, c->stream));
  k_ppm_measure<<<n_cta, B200_PPM_CTA, 0, c->stream>>>(d_sum, npix, (double)total_spp, d_rgb, d_cta);
  k_ppm_scan<<<1, B200_PPM_CTA, 0, c->stream>>>(d_cta, n_cta, (unsigned long long)hl, d_total);
  k_ppm_write<<<n_cta, B200_PPM_CTA, 0, c->stream>>>(d_rgb, npix, d_cta, d_out, dev_cap);
  CU(cudaGetLastError());
  unsigned long long total = 0;
  CU(cudaMemcpyAsync(&total, d_total, 8, cudaMemcpyDeviceToHost, c->stream));
  CU(cudaStreamSynchronize(c->stream));
  *out_len = (size_t)total;
  if (total > capacity) return fail(B200RT_ERR_INVALID, "format_ppm: the file needs %llu bytes, capacity is %zu (b200rt_ppm_bound() always suffices)", total, capacity);
  CU(cudaMemcpyAsync(out + hl, d_out + hl, (size_t)total - hl, cudaMemcpyDeviceToHost, c->stream));
  CU(cudaStreamSynchronize(c->stream));
  std::memcpy(out, header, hl);
  return B200RT_OK;
}

} // extern "C"

// ------------------------------------------------------------------------------------------------ binary image writers
namespace {
void exr_attr(std::string &h, const char *name, const char *type, const void *data, int32_t size) {
  h.append(name).push_back('\0');
  h.append(type).push_back('\0');
  h.append(reinterpret_cast<const char *>(&size), 4);
  h.append(reinterpret_cast<const char *>(data), (size_t)size);
}
// magic, version, attributes, terminator; padded (with a trailing comment attribute) so that header + offset table end on a multiple of 4
std::string exr_header(int W, int H) {
  std::string h;
  const unsigned char magic[8] = {0x76, 0x2f, 0x31, 0x01, 2, 0, 0, 0};
  h.append(reinterpret_cast<const char *>(magic), 8);
  std::string ch;
  for (const char *nm : {"B", "G", "R"}) {
    ch.append(nm).push_back('\0');
    const int32_t type = 2 /* FLOAT */, samp = 1;
    const unsigned char lin[4] = {0, 0, 0, 0};
    ch.append(reinterpret_cast<const char *>(&type), 4);
    ch.append(reinterpret_cast<const char *>(lin), 4);
    ch.append(reinterpret_cast<const char *>(&samp), 4);
    ch.append(reinterpret_cast<const char *>(&samp), 4);
  }
  ch.push_back('\0');
  exr_attr(h, "channels", "chlist", ch.data(), (int32_t)ch.size());
  const unsigned char none = 0;
  exr_attr(h, "compression", "compression", &none, 1);
  const int32_t win[4] = {0, 0, W - 1, H - 1};
  exr_attr(h, "dataWindow", "box2i", win, 16);
  exr_attr(h, "displayWindow", "box2i", win, 16);
  exr_attr(h, "lineOrder", "lineOrder", &none, 1);
  const float one = 1.0f, zero2[2] = {0.0f, 0.0f};
  exr_attr(h, "pixelAspectRatio", "float", &one, 4);
  exr_attr(h, "screenWindowCenter", "v2f", zero2, 8);
  exr_attr(h, "screenWindowWidth", "float", &one, 4);
  const std::string base = "b200rt linear radiance (pixel_color / samples_per_pixel), no gamma";
  // "comments" string attribute sized so that the scan-line blocks are 4-byte aligned
  const size_t fixed = h.size() + 9 + 7 + 4 + 1; // name, type, size field, header terminator
  std::string text = base;
  while ((fixed + text.size()) % 4 != 0) text.push_back(' ');
  exr_attr(h, "comments", "string", text.data(), (int32_t)text.size());
  h.push_back('\0');
  return h;
}
} // namespace

extern "C" {

size_t b200rt_image_bound(int format, int image_width, int image_height) {
  if (image_width <= 0 || image_height <= 0) return 0;
  const size_t W = (size_t)image_width, H = (size_t)image_height;
  switch (format) {
  case B200RT_IMAGE_PPM_TEXT: return b200rt_ppm_bound(image_width, image_height);
  case B200RT_IMAGE_PPM_BINARY: return 64 + W * H * 3;
  case B200RT_IMAGE_EXR: return exr_header(image_width, image_height).size() + 8 * H + H * (8 + 12 * W);
  }
  return 0;
}

int b200rt_format_image(b200rt_ctx *c, const double *rgb_sum, int total_spp, int format, char *out, size_t capacity, size_t *out_len) {
  if (format == B200RT_IMAGE_PPM_TEXT) return b200rt_format_ppm(c, rgb_sum, total_spp, out, capacity, out_len);
  if (!c || !rgb_sum || !out || !out_len) return fail(B200RT_ERR_INVALID, "NULL argument");
  if (!c->have_camera) return fail(B200RT_ERR_STATE, "format_image before b200rt_set_camera");
  if (total_spp <= 0) return fail(B200RT_ERR_INVALID, "total_spp must be positive");
  if (format != B200RT_IMAGE_PPM_BINARY && format != B200RT_IMAGE_EXR) return fail(B200RT_ERR_INVALID, "unknown image format %d", format);
  const int W = c->cam.image_width, H = c->cam.image_height;
  const size_t npix = (size_t)W * H, need = b200rt_image_bound(format, W, H);
  std::string head;
  if (format == B200RT_IMAGE_PPM_BINARY) {
    char header[64];
    head.assign(header, (size_t)snprintf(header, sizeof header, "P6\n%d %d\n255\n", W, H));
  } else {
    head = exr_header(W, H);
    const uint64_t first = head.size() + 8ull * (uint64_t)H;
    for (int y = 0; y < H; y++) {
      const uint64_t off = first + (uint64_t)y * (8ull + 12ull * (uint64_t)W);
      head.append(reinterpret_cast<const char *>(&off), 8);
    }
  }
  const size_t payload = format == B200RT_IMAGE_PPM_BINARY ? npix * 3 : (size_t)H * (8 + 12 * (size_t)W);
  *out_len = head.size() + payload;
  if (*out_len > capacity) return fail(B200RT_ERR_INVALID, "format_image: the file needs %zu bytes, capacity is %zu (b200rt_image_bound() gives %zu)", *out_len, capacity, need);
  CU(cudaSetDevice(c->device));
  DevMem tmp;
  double *d_sum = nullptr;
  unsigned char *d_out = nullptr;
  CU(tmp.alloc((void **)&d_sum, npix * 3 * sizeof(double)));
  CU(tmp.alloc((void **)&d_out, payload));
  CU(cudaMemcpyAsync(d_sum, rgb_sum, npix * 3 * sizeof(double), cudaMemcpyHostToDevice, c->stream));
  if (format == B200RT_IMAGE_PPM_BINARY)
    k_pack_p6<<<grid_for(c, npix * 3, 256, 8), 256, 0, c->stream>>>(d_sum, npix * 3, (double)total_spp, d_out);
  else
    k_pack_exr<<<grid_for(c, npix, 256, 8), 256, 0, c->stream>>>(d_sum, W, H, (double)total_spp, d_out);
  CU(cudaGetLastError());
  CU(cudaMemcpyAsync(out + head.size(), d_out, payload, cudaMemcpyDeviceToHost, c->stream));
  CU(cudaStreamSynchronize(c->stream));
  std::memcpy(out, head.data(), head.size());
  return B200RT_OK;
}

} // extern "C"

#include "group.cuh"
#include "noalloc.cuh"
