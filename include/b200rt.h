/*
 * b200rt.h — C ABI of libb200rt.so, the B200 (sm_100a) path-tracing backend that
 * replaces the per-pixel hot path of JTtNinjaCode/CPU-Ray-Tracing-Implementation:
 *
 *     camera::render's sample loop      (reference src/camera.h:135-176)
 *     camera::ray_color  (recursive)    (reference src/camera.h:193-241)
 *     camera::generate_ray / miss       (reference src/camera.h:244-284, 180-190)
 *     and everything they call per sample (hittable::hit, material::scatter,
 *     pdf::value/generate, texture::sample).
 *
 * The reference has no FFI: the path sits behind a C++ member call.  The C++
 * adaptor in include/b200rt_adaptor.hpp walks the reference's polymorphic scene
 * graph and fills the plain-old-data description below; this header is what
 * that adaptor (and the ctypes mirror in cpu-ray-tracing-implementation_b200/)
 * binds.  No C++ or torch types cross this boundary: plain pointers and sizes.
 *
 * Conventions
 *   - every entry point returns 0 (B200RT_OK) or a negative b200rt_status;
 *     b200rt_last_error() gives the message of the calling thread's last failure.
 *   - the caller owns every host buffer; the library copies during the call.
 *   - one context per GPU; contexts are independent; a context is not re-entrant.
 *   - all geometry and radiance arithmetic is IEEE binary64, in the reference's
 *     operation order and without fused multiply-add, so hit results are
 *     bit-identical to the reference's x86-64 build.
 */
#ifndef B200RT_H
#define B200RT_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200RT_ABI_VERSION 2 /* 2: b200rt_scene_desc::tri_uv, float images, B200RT_TEX_PICTURE_LINEAR */

typedef enum b200rt_status {
  B200RT_OK = 0,
  B200RT_ERR_INVALID = -1,     /* bad argument / malformed scene description   */
  B200RT_ERR_CUDA = -2,        /* a CUDA runtime call failed                    */
  B200RT_ERR_UNSUPPORTED = -3, /* scene uses a construct the backend rejects    */
  B200RT_ERR_STATE = -4        /* call order (render before upload, ...)        */
} b200rt_status;

/* ---------------------------------------------------------------- scene graph
 * A node-for-node mirror of the reference's hittable graph, expanded to a tree
 * (an object reachable along two paths is emitted twice, except the
 * left_ == right_ case of one-object bvh_node, reference src/bvh_node.h:34-36,
 * which keeps a == b so that the double visit is reproduced).
 */
typedef enum b200rt_node_kind {
  B200RT_NODE_BVH = 0,       /* bvh_node       src/bvh_node.h:49-59   a=left b=right f=box{x0,x1,y0,y1,z0,z1} */
  B200RT_NODE_LIST = 1,      /* hittable_list  src/hittable_list.h:20-31  a=first index in list_children, b=count */
  B200RT_NODE_TRANSLATE = 2, /* translate      src/hittable.h:75-82   a=child f[0..2]=offset */
  B200RT_NODE_ROTATE_X = 3,  /* rotate_x       src/hittable.h:125-149 a=child f[0]=sin f[1]=cos */
  B200RT_NODE_ROTATE_Y = 4,  /* rotate_y       src/hittable.h:192-216 */
  B200RT_NODE_ROTATE_Z = 5,  /* rotate_z       src/hittable.h:259-284 */
  B200RT_NODE_SPHERE = 6,    /* sphere         src/sphere.h:40-74     a=index into spheres, c=primitive id */
  B200RT_NODE_QUAD = 7,      /* quad           src/quad.h:30-52       a=index into quads,   c=primitive id */
  B200RT_NODE_TRIANGLE = 8,  /* triangle       src/triangle.h:27-40   a=index into tris,    c=primitive id */
  B200RT_NODE_VOLUME = 9     /* volumne        src/volumne.h:18-46    a=index into volumes, c=primitive id */
} b200rt_node_kind;

typedef struct b200rt_node {
  int32_t kind; /* b200rt_node_kind */
  int32_t a;
  int32_t b;
  int32_t c; /* leaves: primitive id (order of first visit in a left-to-right depth-first
                walk from the root; what b200rt_trace returns); -1 inside a volume boundary */
  double f[6];
} b200rt_node;

/* Geometry records, binary64, one row per object. */
#define B200RT_SPHERE_STRIDE 12 /* center_[3] radius is_moving begin_center_[3] end_center_[3] pad  (src/sphere.h:98-106) */
#define B200RT_QUAD_STRIDE 20   /* corner_[3] u_[3] v_[3] unorm_[3] w[3] D area pad[3]              (src/quad.h:81-87);
                                   w = n/dot(n,n) with n = cross(u_,v_) and D = dot(unorm_,corner_) are
                                   what quad::hit recomputes on every call (src/quad.h:32,39-40)          */
#define B200RT_TRI_STRIDE 12    /* p0_[3] p1_[3] p2_[3] normal_[3]                                  (src/triangle.h:51-55) */

typedef struct b200rt_volume {
  int32_t boundary_node; /* root of the boundary sub-graph (leaves there carry c = -1) */
  int32_t material;      /* the isotropic phase function, src/volumne.h:14 */
  double density;
} b200rt_volume;

typedef enum b200rt_material_kind {
  B200RT_MAT_LAMBERTIAN = 0,   /* src/material.h:57-76   */
  B200RT_MAT_METAL = 1,        /* src/material.h:78-97   p0 = fuzz_ (float)  */
  B200RT_MAT_DIELECTRIC = 2,   /* src/material.h:100-143 p0 = refract_ (float) */
  B200RT_MAT_GLOSS = 3,        /* src/material.h:145-185 p0 = smoothness_ p1 = specular_prob_ (floats) */
  B200RT_MAT_ISOTROPIC = 4,    /* src/material.h:187-204 */
  B200RT_MAT_DIFFUSE_LIGHT = 5 /* src/material.h:206-219 */
} b200rt_material_kind;

typedef struct b200rt_material {
  int32_t kind;
  int32_t tex;
  float p0;
  float p1;
} b200rt_material;

typedef enum b200rt_texture_kind {
  B200RT_TEX_SOLID = 0,   /* src/texture.h:12-37   c0 = colour */
  B200RT_TEX_CHECKER = 1, /* src/texture.h:39-62   c0 = odd c1 = even scale */
  B200RT_TEX_PICTURE = 2, /* src/texture.h:64-78   aux = image index */
  B200RT_TEX_PERLIN = 3,  /* src/texture.h:80-93   aux = perlin table index, scale */
  B200RT_TEX_VALUE = 4,   /* src/texture.h:95-103  aux = value-noise index */
  B200RT_TEX_WORLEY = 5,  /* src/texture.h:105-111 */
  B200RT_TEX_VORONOI = 6, /* src/texture.h:113-119 */
  B200RT_TEX_PICTURE_LINEAR = 7 /* NOT in the reference (opt-in): picture_texture with bilinear filtering between
                                   texel centres, edges clamped; aux = image index.  The reference's lookup is nearest (src/texture.h:67-74) */
} b200rt_texture_kind;

typedef struct b200rt_texture {
  int32_t kind;
  int32_t aux;
  double c0[3];
  double c1[3];
  double scale;
} b200rt_texture;

typedef struct b200rt_image { /* src/image.h:85-90; bytes are image::bdata, already 8-bit linear */
  int32_t width;
  int32_t height;
  int32_t bytes_per_pixel; /* 3 (stb) or 4 (EXR): the reference's 8-bit image::bdata, read as byte / 256 (src/texture.h:72-73).
                              12 or 16: 3 or 4 binary32 channels per pixel = image::fdata, un-quantised and unclamped (opt-in:
                              the reference always quantises, src/image.h:99-117); offset must then be a multiple of 4 */
  int32_t has_data;        /* 0: bdata == nullptr, lookups give magenta (src/image.h:75-76) */
  uint64_t offset;         /* into image_bytes */
} b200rt_image;

typedef struct b200rt_value_noise { /* src/noise.h:95-137 */
  int32_t resolution;
  int32_t pad;
  uint64_t offset; /* into value_lattice (floats), resolution^3 entries */
} b200rt_value_noise;

typedef enum b200rt_light_kind {
  B200RT_LIGHT_NONE = 0,   /* render(out, world)           : material pdf only, src/camera.h:217-226 */
  B200RT_LIGHT_QUAD = 1,   /* quad::pdf_value / random     : src/quad.h:66-78                          */
  B200RT_LIGHT_SPHERE = 2, /* sphere::pdf_value / random   : src/sphere.h:76-81                        */
  B200RT_LIGHT_DEFAULT = 3 /* hittable base: pdf 0, direction (1,0,0): src/hittable.h:39-41           */
} b200rt_light_kind;

typedef struct b200rt_scene_desc {
  uint32_t abi_version; /* B200RT_ABI_VERSION */
  uint32_t root;        /* node index of the world passed to camera::render */

  uint32_t n_nodes;
  uint32_t n_list_children;
  const b200rt_node *nodes;
  const uint32_t *list_children;

  uint32_t n_spheres;
  uint32_t n_quads;
  uint32_t n_tris;
  uint32_t n_volumes;
  const double *spheres;      /* n_spheres x B200RT_SPHERE_STRIDE */
  const int32_t *sphere_mat;  /* material index or -1 (nullptr material) */
  const double *quads;        /* n_quads x B200RT_QUAD_STRIDE */
  const int32_t *quad_mat;
  const double *tris;         /* n_tris x B200RT_TRI_STRIDE */
  const int32_t *tri_mat;
  const b200rt_volume *volumes;

  uint32_t n_materials;
  uint32_t n_textures;
  const b200rt_material *materials;
  const b200rt_texture *textures;

  uint32_t n_images;
  uint32_t n_perlin;
  uint32_t n_value_noise;
  uint32_t n_primitives; /* number of distinct primitive ids (max c + 1) */
  const b200rt_image *images;
  const uint8_t *image_bytes;
  uint64_t image_bytes_size;
  const double *perlin_vec;   /* n_perlin x 256 x 3 : perlin::rand_offset (src/noise.h:72) */
  const int32_t *perlin_perm; /* n_perlin x 256     : perlin::perm_x      (src/noise.h:73; the only table read, :36) */
  const b200rt_value_noise *value_noise;
  const float *value_lattice;
  uint64_t value_lattice_size;

  int32_t light_kind;   /* b200rt_light_kind: the `light` argument of camera::render (src/camera.h:135) */
  int32_t light_index;  /* row in quads / spheres */
  int32_t background_tex; /* camera::background_ (src/camera.h:329); -1 = nullptr */
  int32_t reserved;
  const double *tri_uv; /* NULL (what the reference does: triangle::hit leaves rec.u, rec.v unwritten, src/triangle.h:35-39; the
                           ABI reports 0) or n_tris x {u0 v0 u1 v1 u2 v2}: texture coordinates of the three vertices, interpolated
                           with the barycentric weights of the hit (opt-in, SURVEY.md 8f rank 3: glTF TEXCOORD_0) */
} b200rt_scene_desc;

/* ------------------------------------------------------------------- camera
 * The public fields of the reference camera after initialize_* (src/camera.h:298-329).
 */
typedef enum b200rt_camera_mode {
  B200RT_CAM_PERSPECTIVE = 0, /* src/camera.h:245-251 */
  B200RT_CAM_ORTHOGRAPHIC = 1, /* src/camera.h:252-258 */
  B200RT_CAM_FISHEYE = 2,     /* src/camera.h:259-275 */
  B200RT_CAM_LENS = 3         /* src/camera.h:276-283 */
} b200rt_camera_mode;

typedef struct b200rt_camera {
  int32_t mode;
  int32_t image_width;
  int32_t image_height;
  int32_t samples_per_pixel;
  int32_t max_depth;
  int32_t reserved;
  double viewport_width;
  double viewport_height;
  double focal_length;
  double focus_dist;
  double pos[3];
  double dir[3];
  double right[3];
  double up[3];
  double defocus_disk_u[3];
  double defocus_disk_v[3];
} b200rt_camera;

/* ------------------------------------------------------------------ options */
typedef enum b200rt_option {
  B200RT_OPT_TRAVERSAL = 0,     /* 0 = compacted SAH BVH (default), 1 = node-for-node mirror of the reference graph */
  B200RT_OPT_WAVE_PATHS = 1,    /* paths resident per wave, shared by the pipelines (default 32 Mi = 10 GB of wavefront state) */
  B200RT_OPT_KILL_ZERO_WEIGHT = 2, /* 1 (default): drop paths whose throughput is exactly 0 in all channels */
  B200RT_OPT_BVH_LEAF_SIZE = 3, /* max primitives per leaf of the compacted BVH; 0 (default) = 1 up to 8192 primitives, else 2 */
  /* 4 was B200RT_OPT_SMEM_NODES (never implemented; removed: b200rt_set_option rejects it) */
  B200RT_OPT_EXTEND_BLOCKS = 6, /* tuning: resident CTAs per SM the persistent extend kernel is compiled for (4 or 6; 6 spills and is slower: profiles/r01_g_tune.txt) */
  B200RT_OPT_SPECULATIVE_LEAF = 7, /* tuning: 1 = a lane postpones its first leaf and keeps walking inner nodes (speculative traversal) */
  B200RT_OPT_FUSE_MISS = 8,     /* 1 (default): in a scene without background the extend kernel finishes a miss itself (it adds 0) instead of queueing the path */
  B200RT_OPT_EXTEND_STAGES = 9, /* tuning: ray batches staged in shared memory per warp, 2 (default) or 1; lane refill always uses 2 */
  B200RT_OPT_SMEM_CARVEOUT = 10, /* tuning: shared-memory carve-out of the extend kernel in percent, -1 (default) = smallest that fits */
  B200RT_OPT_BVH_BUILDER = 11,  /* who builds the compacted BVH that replaces bvh_node's constructor (src/bvh_node.h:18-47): B200RT_BVH_HOST_SAH (default), B200RT_BVH_DEVICE_LBVH or B200RT_BVH_DEVICE_PLOC */
  B200RT_OPT_BVH_WIDTH = 12,    /* children per node of the traversed tree: 2, 4 (collapsed from the host SAH BVH2), 0 (default) = 4 above 2048 primitives */
  B200RT_OPT_PIPELINES = 13,    /* 1 or 2 (default): with 2 the sample range is split into two independent wavefronts on two streams; each half's kernels fill the SMs the other half's kernel has drained */
  B200RT_OPT_PIPE_EXTEND_CTAS = 14, /* tuning: resident extend CTAs per SM in two-pipeline mode (1..4, default 4; fewer leave registers for the other half's kernels and were measured slower) */
  B200RT_OPT_TAIL_PATHS = 15, /* tuning: once every sample is generated and at most this many paths are alive, one kernel follows each of them to its
                                 end instead of one wave per bounce (default 32768; 0 = off).  Same draws, same contributions */
  B200RT_OPT_WAVE_LOOP = 16,  /* who drives the wave loop of a render call: 1 (default) = the device: one CUDA graph per pipeline whose while node repeats
                                 the wave until no path is alive and no sample is left (no host round trip; the call returns without waiting unless
                                 statistics are requested); 0 = the host, reading the counters back (the mode that can time single launches:
                                 b200rt_stats::extend_ms).  The mirror traversal always uses the host loop */
  B200RT_OPT_EXTEND_LOOP = 17, /* tuning: loop structure of the persistent extend kernel: 0 = while-while (inner nodes until every lane of the warp has
                                  found a leaf, then the leaves), 1 = warp-uniform "threshold if-if": the leaf phase starts as soon as
                                  B200RT_OPT_LEAF_LANES lanes wait at a leaf; -1 (default) = by scene: 1 with 20 lanes for a triangle mesh behind
                                  the BVH4 (>= 90 % of the leaves are triangles), 0 otherwise */
  B200RT_OPT_LEAF_LANES = 18,  /* tuning: see B200RT_OPT_EXTEND_LOOP (1..32; 0 = default: 20 when the scene chose the loop, 8 when the caller did;
                                  32 = the while-while schedule) */
  B200RT_OPT_L2_PERSIST = 19,  /* 1 (default): the tree, the leaf entries and the primitive rows are uploaded into one address range that the render
                                  kernels declare as a persisting L2 access-policy window (the wave's path state streams past it) */
  B200RT_OPT_TRI_F32 = 20,     /* 1 (default): triangle vertices are kept as 48-byte binary32 rows when every coordinate is exactly representable
                                  (glTF positions are), and widened after the load: identical binary64 operands, 40 % fewer bytes per test */
  B200RT_OPT_OCTANT_SORT = 21, /* tuning: 1 = the shade kernels write the surviving paths of a CTA iteration grouped by the octant of their new
                                  direction (more coherent 32-ray batches for the extend kernel; same paths, same samples) */
  B200RT_OPT_RAY_SORT = 22,    /* experiment: 0 (default) = off; else the live paths of every wave are radix-sorted by a key of their ray before the
                                  extend kernel (bit 0: direction octant, bit 1: 16 cells per face of the direction cube, bit 2: origin cell on a
                                  16^3 grid) and the extend kernel gathers its 32-ray batches in that order.  Same paths, same results */
  B200RT_OPT_REFILL_LANES = 5   /* idle lanes per warp at which the persistent extend kernel claims new rays (32 = fixed 32-ray batches; 0 (default) = 32, or 8 from 32 Ki BVH nodes) */
} b200rt_option;

#define B200RT_BVH_HOST_SAH 0    /* binned SAH over three axes on the host: best tree, ~0.5 us per primitive */
#define B200RT_BVH_DEVICE_LBVH 1 /* Morton-order linear BVH built by CUDA kernels: one primitive per leaf, lower quality, ~1000x faster to build */
#define B200RT_BVH_DEVICE_PLOC 2 /* topology by parallel locally-ordered clustering on the device (csrc/bvh_ploc.cuh), leaf pairs + layout + BVH4 collapse on the host: close to the SAH tree's traversal speed at a fraction of its build time */
#define B200RT_TRAVERSAL_FAST 0
#define B200RT_TRAVERSAL_MIRROR 1

typedef struct b200rt_stats {
  uint64_t samples;       /* iterations of the k loop, src/camera.h:165-168             */
  uint64_t rays;          /* world.hit calls issued from ray_color, src/camera.h:198    */
  uint64_t nan_samples;   /* samples whose camera ray had a NaN direction (fisheye, outside the image circle) */
  uint64_t kernel_launches;
  double gpu_ms;          /* CUDA-event time of the kernels of this call                */
  double extend_ms;       /* of which: traversal kernels                                 */
  double h2d_bytes;
  double d2h_bytes;
  uint64_t waves;         /* wavefront iterations = launches of the extend kernel         */
} b200rt_stats;

typedef struct b200rt_ctx b200rt_ctx;

/* Create a context on CUDA device `device`. */
int b200rt_create(int device, b200rt_ctx **out);
int b200rt_destroy(b200rt_ctx *ctx);
const char *b200rt_last_error(void);
int b200rt_abi_version(void);

int b200rt_set_option(b200rt_ctx *ctx, int option, double value);

/* Replaces: scene hand-over at camera::render(out, world, light) (src/camera.h:135) plus the
 * bvh_node constructor (src/bvh_node.h:18-47): copies the description, builds the compacted
 * BVH on the host and uploads everything once. */
int b200rt_upload_scene(b200rt_ctx *ctx, const b200rt_scene_desc *scene);

/* Replaces: camera::initialize_* hand-over (src/camera.h:21,52,74,104). */
int b200rt_set_camera(b200rt_ctx *ctx, const b200rt_camera *camera);

/* Replaces: world.hit(r, interval(tmin,tmax), record) (src/camera.h:198 -> src/bvh_node.h:49-59)
 * on a batch.  rays7 = n x {origin[3], direction[3], time}.  Host buffers.  Any output may be NULL.
 *   prim  : primitive id of the returned record, -1 on miss
 *   t     : record.t                          uv   : record.u, record.v (n x 2)
 *   front : record.front_face                 p    : record.p (n x 3)     normal: record.normal (n x 3)
 * `seed` keys the free-flight draw of volumes (src/volumne.h:36); ray i uses stream i. */
int b200rt_trace(b200rt_ctx *ctx, const double *rays7, size_t n, double tmin, double tmax, uint64_t seed,
                 int32_t *prim, double *t, double *uv, uint8_t *front, double *p, double *normal);

/* Replaces: generate_ray(i, j, du, dv) (src/camera.h:244-284) for samples [spp_begin, spp_end) of every
 * pixel; writes rays7 for pixel-major order  index = (sample - spp_begin) * W*H + (i*W + j).  Host buffer. */
int b200rt_generate_rays(b200rt_ctx *ctx, uint64_t seed, int spp_begin, int spp_end, double *rays7);

/* Replaces: the sample loop of camera::render (src/camera.h:158-172) for samples [spp_begin, spp_end):
 * rgb_sum[(i*W+j)*3 + c] receives the SUM of ray_color over that sample range (divide by the total
 * sample count after combining ranges).  Host buffer; the device->host copy is part of the call. */
int b200rt_render(b200rt_ctx *ctx, uint64_t seed, int spp_begin, int spp_end, double *rgb_sum, b200rt_stats *stats);

/* Same, accumulating (+=) into a DEVICE buffer of W*H*3 doubles with every kernel enqueued on `cuda_stream`
 * (a cudaStream_t, may be 0).  The wave loop is driven from the host, so the call returns once the last wave
 * has drained; the buffer is then complete on that stream.  This is what the multi-GPU driver reduces with NCCL. */
int b200rt_render_device(b200rt_ctx *ctx, uint64_t seed, int spp_begin, int spp_end, double *d_rgb_sum,
                         void *cuda_stream, b200rt_stats *stats);

/* Replaces: pixel_color /= spp and write_color (src/camera.h:169, src/color.h:16-36):
 * rgb_linear = rgb_sum / total_spp (may alias); ppm_rgb (may be NULL) gets int(255.999 * pow(c, 1/2.2)), unclamped. */
int b200rt_finalize(b200rt_ctx *ctx, const double *rgb_sum, int total_spp, double *rgb_linear, int32_t *ppm_rgb);

/* Replaces the output stage of camera::render: the "P3\nW H\n255\n" header (src/camera.h:149-151) and the
 * std::for_each(image_, write_color) loop (src/camera.h:174, src/color.h:22-36).  `out` receives the complete file image
 * exactly as the reference's ofstream would hold it (one "r g b\n" line per pixel, integers unclamped, no terminator);
 * *out_len = its length.  The integer-to-text conversion runs on the device.  A `capacity` of
 * b200rt_ppm_bound(W, H) always suffices; a smaller one fails with B200RT_ERR_INVALID and *out_len = the size needed. */
size_t b200rt_ppm_bound(int image_width, int image_height);
int b200rt_format_ppm(b200rt_ctx *ctx, const double *rgb_sum, int total_spp, char *out, size_t capacity, size_t *out_len);

/* The same output stage in other containers ("binary PPM/EXR writer beside the reference's P3 text").
 *   B200RT_IMAGE_PPM_TEXT   = b200rt_format_ppm: the reference's own file (src/camera.h:149-151, 174)
 *   B200RT_IMAGE_PPM_BINARY   "P6\nW H\n255\n" + one byte per component: the integer write_color prints (src/color.h:22-36), clamped to [0, 255]
 *   B200RT_IMAGE_EXR          OpenEXR, single part, scan lines, no compression, 32-bit float B G R: the LINEAR image pixel_color / spp
 *                             (src/camera.h:169), no gamma, no clamp (emitters and NaNs survive)
 * Pixel conversion and file layout run on the device; `out` receives the complete file image; b200rt_image_bound() always suffices. */
typedef enum b200rt_image_format { B200RT_IMAGE_PPM_TEXT = 0, B200RT_IMAGE_PPM_BINARY = 1, B200RT_IMAGE_EXR = 2 } b200rt_image_format;
size_t b200rt_image_bound(int format, int image_width, int image_height);
int b200rt_format_image(b200rt_ctx *ctx, const double *rgb_sum, int total_spp, int format, char *out, size_t capacity, size_t *out_len);

/* No counterpart in the reference: added for callers that must not reach cudaMalloc (code under
 * CUDA graph capture, tight serving loops, harnesses that forbid it).  The same P3 file image as b200rt_format_ppm, produced by the
 * same kernels, but this entry point allocates nothing: no cudaMalloc / cudaFree / allocating library routine on any path.
 *   d_rgb_sum : DEVICE pointer, W*H*3 doubles (e.g. what b200rt_render_device accumulated); W, H = the context's camera
 *   d_out     : DEVICE pointer, `capacity` bytes; receives the complete file image, header included
 *   *out_len  : host; the file's length (also on B200RT_ERR_INVALID for a too small capacity: the size needed)
 * The scratch of the two-pass formatter (three integers per pixel, per-CTA byte counts, the total: 12 bytes per pixel + 8.25 KiB) lives in
 * a file-scope __device__ array sized for B200RT_NOALLOC_MAX_PIXELS pixels (12 MiB + 8.25 KiB, held by every GPU that loads the library
 * until the process exits).  An image with more pixels fails with B200RT_ERR_UNSUPPORTED.  Every call on a device shares that array, and
 * it keeps the previous call's contents: each call overwrites every word it later reads (nothing has to start at zero), but two calls
 * that overlap on the same GPU (two contexts, two threads) race on it: serialise them.  Work is issued on the context's stream;
 * the call returns after that stream has drained. */
#define B200RT_NOALLOC_MAX_PIXELS (1u << 20)
int b200rt_format_ppm_noalloc(b200rt_ctx *ctx, const double *d_rgb_sum, int total_spp, char *d_out, size_t capacity, size_t *out_len);

/* Replaces: texture::sample(u, v, p) (src/texture.h:9) of texture row `tex` on a batch.
 * uvp5 = n x {u, v, p[3]}, rgb = n x 3.  Host buffers. */
int b200rt_texture_sample(b200rt_ctx *ctx, int tex, const double *uvp5, size_t n, double *rgb);

/* Replaces: one level of camera::ray_color below the hit test (src/camera.h:202-240) on a batch of hit records:
 * material::emitted and material::scatter (src/material.h:57-219), then for scatter_mode::kRandom the pdf the reference
 * builds (scatter_record.pdf alone, or dual_pdf(hittable_pdf(light), scatter_record.pdf), src/camera.h:217-231;
 * src/pdf.h:15-61; src/hittable_list.h:39-50; src/quad.h:66-78; src/sphere.h:76-81), its generate() and value(), and
 * material::p_scattered.  Parity entry: the random numbers are SUPPLIED, so that the result can be compared bit for bit with
 * the reference run on the same std::rand values.  Host buffers.
 *   rays7  : n x {origin[3], direction[3], time}                     the incoming ray r_in
 *   rec12  : n x {p[3], normal[3], t, u, v, front_face, material row, 0}   the hit_record (src/hittable.h:7-30)
 *   draws6 : n x {m0, m1, m2, s0, s1, s2} in [0,1):
 *              m* = what scatter() consumes   metal: (m1, m2) = (u1, u2) of random_in_unit_sphere (src/utility.h:30-42)
 *                                             dielectric: m0 = the reflectance test (src/material.h:123)
 *                                             gloss: m0 = the specular test (:161), (m1, m2) = (r1, r2) of random_cosine_direction
 *              s* = what the pdf consumes     s0 = dual_pdf::generate's choice (src/pdf.h:53; unused without a light),
 *                                             (s1, s2) = (r1, r2) of random_cosine_direction, (u1, u2) of random_in_unit_sphere,
 *                                             or the factors of u_ and v_ in quad::random (src/quad.h:76)
 *   out16  : n x {mode, emitted[3], attenuation[3], direction[3], pdf_value, p_scattered, weight[3], 0}
 *              mode 0 = scatter() returned false, 1 = kDetermined, 2 = kRandom; direction = of the scattered ray;
 *              weight = the factor on the incoming radiance: attenuation, or attenuation * p_scattered / pdf_value */
int b200rt_scatter(b200rt_ctx *ctx, const double *rays7, const double *rec12, const double *draws6, size_t n, double *out16);

/* ------------------------------------------------------------------ several GPUs of one box
 * Replaces: the parallel row loop of camera::render (src/camera.h:158) at the scale of a box: the frame is split by SAMPLE
 * RANGE across the devices of a group (samples are independent, src/camera.h:163-170; the scene is replicated), device g of G
 * renders the global sample indices [begin + g*span/G, ...) of every pixel, and the fp64 accumulators are summed onto the
 * first device with one NCCL reduce over NVLink (ncclCommInitAll: single process, one communicator per device) - or by a
 * peer-access kernel.  Philox draws are keyed by the global sample index, so the set of samples - and, up to fp64 summation
 * order, the image - does not depend on the number of devices.  One host thread per device inside every call. */
typedef struct b200rt_group b200rt_group;
#define B200RT_GROUP_OPT_REDUCE 1000 /* b200rt_group_set_option: how the accumulators are combined */
#define B200RT_GROUP_REDUCE_NCCL 0   /* ncclReduce(sum, root = first device) (default) */
#define B200RT_GROUP_REDUCE_PEER 1   /* one kernel on the first device adds the others' buffers through peer (NVLink) loads, in device order */
int b200rt_group_create(const int *devices, int n, b200rt_group **out);
int b200rt_group_destroy(b200rt_group *group);
int b200rt_group_size(const b200rt_group *group);
/* The context of member i (borrowed; member 0 holds the combined image and serves b200rt_finalize / b200rt_format_ppm). */
int b200rt_group_context(b200rt_group *group, int i, b200rt_ctx **out);
/* Any b200rt_option (applied to every member) or B200RT_GROUP_OPT_REDUCE. */
int b200rt_group_set_option(b200rt_group *group, int option, double value);
int b200rt_group_upload_scene(b200rt_group *group, const b200rt_scene_desc *scene);
int b200rt_group_set_camera(b200rt_group *group, const b200rt_camera *camera);
/* b200rt_render over the whole group: rgb_sum (host, W*H*3) receives the sum over samples [spp_begin, spp_end).
 * stats: samples / rays / launches summed over the devices, gpu_ms = the slowest device. */
int b200rt_group_render(b200rt_group *group, uint64_t seed, int spp_begin, int spp_end, double *rgb_sum, b200rt_stats *stats);
const char *b200rt_group_reduce_backend(void); /* e.g. "NCCL 2.27.3 (dlopen)" once a group of more than one device exists */

/* sizeof() of the ABI structures as compiled into the library (0 scene_desc, 1 camera, 2 stats, 3 node, 4 material,
 * 5 texture, 6 image, 7 volume, 8 value_noise, 9 bvh_info): lets foreign-language bindings check their mirrors. */
size_t b200rt_abi_sizeof(int which);

/* Introspection used by tests and the benchmark. */
typedef struct b200rt_bvh_info {
  uint32_t n_nodes;
  uint32_t n_leaves;
  uint32_t n_prims;
  uint32_t max_depth;
  double build_ms;
  double sah_cost;
} b200rt_bvh_info;
int b200rt_bvh_info_get(b200rt_ctx *ctx, b200rt_bvh_info *out);

#ifdef __cplusplus
}
#endif
#endif /* B200RT_H */
