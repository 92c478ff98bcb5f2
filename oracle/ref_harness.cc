// ref_harness.cc — TEST INFRASTRUCTURE.  One translation unit that compiles the UNMODIFIED
// reference sources where they lie (/root/reference/src, read-only) into oracle/_ref/libref_harness.so.
// Nothing here is part of the product path; only tests/, __graft_entry__.smoke() and the
// cpu_baseline / --impl reference legs of bench.py may load it.
//
// What it gives the tests (all through the reference's own code):
//   * the scenes of BASELINE.json (C1..C5) and of the reference's menu (src/main.cc:634-657),
//     built with the reference's constructors; parameters restate src/main.cc, cited per scene
//   * the flattened description of each scene (include/b200rt_adaptor.hpp walks the live objects)
//   * camera::generate_ray, <world>.hit (= bvh_node::hit / hittable_list::hit), camera::ray_color,
//     texture::sample on caller-supplied batches -> golden vectors for tests/golden/
//   * per-pixel sum / sum-of-squares renders and timings for the CPU baseline
//
// Build: oracle/Makefile (g++ -O2 -std=c++17 -fpermissive -ffp-contract=off, no -march: SURVEY.md §8c).
#include <algorithm>
#include <any>
#include <atomic>
#include <chrono>
#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <execution>
#include <fstream>
#include <functional>
#include <initializer_list>
#include <iostream>
#include <limits>
#include <map>
#include <memory>
#include <mutex>
#include <numeric>
#include <sstream>
#include <stdexcept>
#include <string>
#include <thread>
#include <tuple>
#include <typeinfo>
#include <unordered_map>
#include <utility>
#include <vector>

#include <dlfcn.h>

#define STB_IMAGE_IMPLEMENTATION
#define STBI_FAILURE_USERMSG
#include "json.h" // nlohmann/json must be parsed before the access macro below (libstdc++ <any> breaks otherwise)

// The reference keeps all scene state private without getters.
#define private public
#include "aabb.h"
#include "bvh_node.h"
#include "camera.h"
#include "color.h"
#include "gltf_loader.h"
#include "hittable.h"
#include "hittable_list.h"
#include "onb.h"
#include "quad.h"
#include "ray.h"
#include "sphere.h"
#include "stb_image.h"
#include "triangle.h"
#include "utility.h"
#include "vec3.h"
#include "volumne.h"
#undef private

// The adaptor's b200::render (the drop-in for camera::render) is compiled here too, so that the test suite can drive the
// whole reference-side binding: reference objects -> adaptor -> C ABI -> GPU -> P3 file.  This checker library must load
// (and serve the oracle tests and the CPU reference arm of bench.py) WITHOUT the product library in the process, so the
// adaptor's b200rt_* calls go through function pointers that ref_render_b200 fills with dlopen / dlsym when it is called.
#include "b200rt.h"
namespace b200dyn {
#define B200DYN(name) decltype(&::name) name = nullptr;
B200DYN(b200rt_create)
B200DYN(b200rt_destroy)
B200DYN(b200rt_last_error)
B200DYN(b200rt_upload_scene)
B200DYN(b200rt_set_camera)
B200DYN(b200rt_render)
B200DYN(b200rt_finalize)
B200DYN(b200rt_ppm_bound)
B200DYN(b200rt_format_ppm)
B200DYN(b200rt_group_create)
B200DYN(b200rt_group_destroy)
B200DYN(b200rt_group_upload_scene)
B200DYN(b200rt_group_set_camera)
B200DYN(b200rt_group_render)
B200DYN(b200rt_group_context)
#undef B200DYN
void *handle = nullptr;
bool bind(const char *path, std::string &err) {
  if (handle) return true;
  handle = dlopen(path, RTLD_NOW | RTLD_LOCAL);
  if (!handle) {
    err = std::string("dlopen ") + path + ": " + dlerror();
    return false;
  }
#define B200DYN(name)                                                                                                  \
  name = (decltype(name))dlsym(handle, #name);                                                                          \
  if (!name) {                                                                                                         \
    err = "libb200rt.so does not export " #name;                                                                       \
    return false;                                                                                                      \
  }
  B200DYN(b200rt_create)
  B200DYN(b200rt_destroy)
  B200DYN(b200rt_last_error)
  B200DYN(b200rt_upload_scene)
  B200DYN(b200rt_set_camera)
  B200DYN(b200rt_render)
  B200DYN(b200rt_finalize)
  B200DYN(b200rt_ppm_bound)
  B200DYN(b200rt_format_ppm)
  B200DYN(b200rt_group_create)
  B200DYN(b200rt_group_destroy)
  B200DYN(b200rt_group_upload_scene)
  B200DYN(b200rt_group_set_camera)
  B200DYN(b200rt_group_render)
  B200DYN(b200rt_group_context)
#undef B200DYN
  return true;
}
} // namespace b200dyn
#define b200rt_create (*b200dyn::b200rt_create)
#define b200rt_destroy (*b200dyn::b200rt_destroy)
#define b200rt_last_error (*b200dyn::b200rt_last_error)
#define b200rt_upload_scene (*b200dyn::b200rt_upload_scene)
#define b200rt_set_camera (*b200dyn::b200rt_set_camera)
#define b200rt_render (*b200dyn::b200rt_render)
#define b200rt_finalize (*b200dyn::b200rt_finalize)
#define b200rt_ppm_bound (*b200dyn::b200rt_ppm_bound)
#define b200rt_format_ppm (*b200dyn::b200rt_format_ppm)
#define b200rt_group_create (*b200dyn::b200rt_group_create)
#define b200rt_group_destroy (*b200dyn::b200rt_group_destroy)
#define b200rt_group_upload_scene (*b200dyn::b200rt_group_upload_scene)
#define b200rt_group_set_camera (*b200dyn::b200rt_group_set_camera)
#define b200rt_group_render (*b200dyn::b200rt_group_render)
#define b200rt_group_context (*b200dyn::b200rt_group_context)
#include "b200rt_adaptor.hpp"

// ------------------------------------------------------------------------------------------------
// std::rand provider.  mode 0: glibc's generator (what the stock executable uses, global + locked,
// src/utility.h:20).  mode 1: a thread-local splitmix/xorshift stream so that rows can be rendered on all
// cores without the glibc lock and with results independent of the thread count.  Linked with
// -Bsymbolic-functions so the reference's calls bind here.
namespace {
int g_rng_mode = 0;
thread_local uint64_t t_rng_state = 0x9E3779B97F4A7C15ull;
typedef int (*rand_fn)(void);
rand_fn libc_rand() {
  static rand_fn f = (rand_fn)dlsym(RTLD_NEXT, "rand");
  return f;
}
inline uint64_t splitmix(uint64_t &s) {
  uint64_t z = (s += 0x9E3779B97F4A7C15ull);
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}
void tls_seed(uint64_t a, uint64_t b) {
  uint64_t s = a * 0xD1342543DE82EF95ull + b;
  splitmix(s);
  t_rng_state = s;
}
} // namespace
thread_local std::vector<int> *t_rand_log = nullptr; // when set, every value handed to the reference is recorded
extern "C" int rand(void) {
  int v = (g_rng_mode == 0) ? libc_rand()() : (int)(splitmix(t_rng_state) >> 33); // 31 bits, 0..RAND_MAX
  if (t_rand_log) t_rand_log->push_back(v);
  return v;
}

// ------------------------------------------------------------------------------------------------
namespace {

using hptr = std::shared_ptr<hittable>;
using mptr = std::shared_ptr<material>;
using tptr = std::shared_ptr<texture>;

struct scene {
  hptr world;
  std::shared_ptr<const hittable> light;
  camera cam;
  b200::flattener flat;
  b200rt_scene_desc desc;
  b200rt_camera cam_desc;
  std::vector<const texture *> tex_by_row;
  std::string name;
};

// counts world.hit calls issued by ray_color (src/camera.h:198); NaN-direction rays are not counted
class counting_world : public hittable {
public:
  explicit counting_world(const hittable &w) : w_(w) {}
  bool hit(const ray &r, interval t, hit_record &rec) const override {
    const vec3 &d = r.direction();
    if (!(std::isnan(d.e[0]) || std::isnan(d.e[1]) || std::isnan(d.e[2]))) counter()++;
    return w_.hit(r, t, rec);
  }
  aabb get_bounding_box() const override { return w_.get_bounding_box(); }
  static uint64_t &counter() {
    static thread_local uint64_t c = 0;
    return c;
  }
  const hittable &w_;
};

// Replace every leaf primitive reachable from `h` by a b200::leaf_tag around it, in place, ids in
// depth-first first-visit order (the order include/b200rt.h documents).  Boxes are untouched, so the
// tree the reference traverses is the tree it built.
struct instrumenter {
  std::unordered_map<const hittable *, hptr> wrapped;
  int next = 0;
  bool is_container(const hittable *h) {
    return dynamic_cast<const bvh_node *>(h) || dynamic_cast<const hittable_list *>(h) || dynamic_cast<const translate *>(h) ||
           dynamic_cast<const rotate_x *>(h) || dynamic_cast<const rotate_y *>(h) || dynamic_cast<const rotate_z *>(h);
  }
  void fix(hptr &slot) {
    hittable *h = slot.get();
    if (dynamic_cast<b200::leaf_tag *>(h)) return;
    if (is_container(h)) {
      descend(h);
      return;
    }
    auto it = wrapped.find(h); // volumne, sphere, quad, triangle: a leaf of the top-level query
    if (it == wrapped.end()) it = wrapped.emplace(h, std::make_shared<b200::leaf_tag>(slot, next++)).first;
    slot = it->second;
  }
  void descend(hittable *h) {
    if (auto n = dynamic_cast<bvh_node *>(h)) {
      bool same = n->left_.get() == n->right_.get();
      fix(n->left_);
      if (same)
        n->right_ = n->left_;
      else
        fix(n->right_);
    } else if (auto l = dynamic_cast<hittable_list *>(h)) {
      for (auto &o : l->objects) fix(o);
    } else if (auto t = dynamic_cast<translate *>(h)) {
      fix(t->object_);
    } else if (auto r = dynamic_cast<rotate_x *>(h)) {
      fix(r->object);
    } else if (auto r = dynamic_cast<rotate_y *>(h)) {
      fix(r->object);
    } else if (auto r = dynamic_cast<rotate_z *>(h)) {
      fix(r->object);
    }
  }
};

// ---------------------------------------------------------------- scene construction helpers
tptr solid(double r, double g, double b) { return std::make_shared<solid_color>(color(r, g, b)); }
tptr solid(double k) { return solid(k, k, k); }
mptr lam(double r, double g, double b) { return std::make_shared<lambertian>(solid(r, g, b)); }
mptr lamt(tptr t) { return std::make_shared<lambertian>(t); }
mptr emit(double k) { return std::make_shared<diffuse_light>(solid(k)); }
hptr S(point3 c, double r, mptr m) { return std::make_shared<sphere>(c, r, m); }
std::shared_ptr<quad> Q(point3 o, vec3 u, vec3 v, mptr m) { return std::make_shared<quad>(o, u, v, m); }
tptr picture(const std::string &path) { return std::make_shared<picture_texture>(std::make_shared<image>(path.c_str())); }

struct build_args {
  int width, spp, depth;
  std::string assets;
  const float *mesh;
  int n_mesh_tris;
};
int pick(int v, int dflt) { return v > 0 ? v : dflt; }

// the five walls + ceiling light shared by the Cornell variants of src/main.cc:227-307
void cornell_shell_b(hittable_list &w, mptr red, mptr white, mptr green) {
  w.push_back(Q(point3(555, 0, 0), vec3(0, 555, 0), vec3(0, 0, 555), green));
  w.push_back(Q(point3(0, 0, 0), vec3(0, 555, 0), vec3(0, 0, 555), red));
  w.push_back(Q(point3(0, 555, 0), vec3(555, 0, 0), vec3(0, 0, 555), white));
  w.push_back(Q(point3(0, 0, 0), vec3(555, 0, 0), vec3(0, 0, 555), white));
  w.push_back(Q(point3(0, 0, 555), vec3(555, 0, 0), vec3(0, 555, 0), white));
}
// the walls of src/main.cc:205-209 (ceiling given from the far corner)
void cornell_shell_a(hittable_list &w, mptr red, mptr white, mptr green) {
  w.push_back(Q(point3(555, 0, 0), vec3(0, 555, 0), vec3(0, 0, 555), green));
  w.push_back(Q(point3(0, 0, 0), vec3(0, 555, 0), vec3(0, 0, 555), red));
  w.push_back(Q(point3(0, 0, 0), vec3(555, 0, 0), vec3(0, 0, 555), white));
  w.push_back(Q(point3(555, 555, 555), vec3(-555, 0, 0), vec3(0, 0, -555), white));
  w.push_back(Q(point3(0, 0, 555), vec3(555, 0, 0), vec3(0, 555, 0), white));
}
// the long gallery of src/main.cc:317-321 / 520-524
void gallery_shell(hittable_list &w) {
  auto red = lam(.65, .05, .05), white = lam(.73, .73, .73), green = lam(.12, .45, .15);
  w.push_back(Q(point3(18, -4, -3), vec3(0, 8, 0), vec3(0, 0, 6), green));
  w.push_back(Q(point3(0, -4, -3), vec3(0, 8, 0), vec3(0, 0, 6), red));
  w.push_back(Q(point3(0, -4, -3), vec3(18, 0, 0), vec3(0, 0, 6), white));
  w.push_back(Q(point3(0, 4, -3), vec3(18, 0, 0), vec3(0, 0, 6), white));
  w.push_back(Q(point3(0, -4, -3), vec3(18, 0, 0), vec3(0, 10, 0), white));
}

std::vector<std::array<vec3, 3>> load_gltf_triangles(const std::string &path) {
  // the copy loops of src/main.cc:351-393: f32 positions, optional u16 indices
  gltf::GltfLoader model(path);
  std::vector<std::array<vec3, 3>> tris;
  for (auto &prim : model.getOutputPrimitives()) {
    std::vector<vec3> pos;
    if (prim.pos_type == DataType::kFloat)
      for (size_t j = 0; j + 12 <= prim.positions.size(); j += 12) {
        float x, y, z;
        memcpy(&x, &prim.positions[j], 4);
        memcpy(&y, &prim.positions[j + 4], 4);
        memcpy(&z, &prim.positions[j + 8], 4);
        pos.push_back(vec3(x, y, z));
      }
    std::vector<unsigned short> idx;
    if (prim.use_indices && prim.indices_type == DataType::kUnsignedShort)
      for (size_t j = 0; j + 2 <= prim.indices.size(); j += 2) {
        unsigned short k;
        memcpy(&k, &prim.indices[j], 2);
        idx.push_back(k);
      }
    if (prim.use_indices)
      for (size_t i = 0; i + 2 < idx.size(); i += 3) tris.push_back({pos[idx[i]], pos[idx[i + 1]], pos[idx[i + 2]]});
    else
      for (size_t i = 0; i + 2 < pos.size(); i += 3) tris.push_back({pos[i], pos[i + 1], pos[i + 2]});
  }
  return tris;
}

void sphereflake_gen(double radius, const vec3 &center, mptr mat, int iteration, hittable_list &world, double scale, vec3 direction) {
  // src/main.cc:23-46
  world.push_back(S(center, radius, mat));
  if (iteration == 0) return;
  onb uvw(direction);
  auto child = [&](vec3 local) {
    vec3 off = uvw.transform(local);
    sphereflake_gen(radius * scale, center + off * (radius + radius * scale), mat, iteration - 1, world, scale, off);
  };
  for (int i = 0; i < 6; ++i) {
    double a = 2.0 * pi * i / 6.0;
    child(vec3(std::cos(a), 0.0, std::sin(a)));
  }
  for (int i = 0; i < 3; ++i) {
    double a = 2.0 * pi * i / 3.0;
    child(vec3(std::cos(a) * std::cos(pi / 3.0), std::sin(pi / 3.0), std::sin(a) * std::cos(pi / 3.0)));
  }
}

// Builds scene `name`.  width/spp/depth <= 0 keep the scene's own values.
void build(scene &sc, const std::string &name, const build_args &a) {
  hittable_list world;
  camera &cam = sc.cam;
  std::shared_ptr<const hittable> light;
  bool use_bvh = false;
  const std::string A = a.assets;

  if (name == "cornell" || name == "cornell_shipped") {
    // C1 (BASELINE.json configs[0]) = geometry of src/main.cc:199-219 at 100 spp / depth 50; "_shipped" = 40 spp / depth 4 (:222)
    auto red = lam(.65, .05, .05), white = lam(.73, .73, .73), green = lam(.12, .45, .15);
    cornell_shell_a(world, red, white, green);
    world.push_back(std::make_shared<translate>(vec3(100, 0, 200), box(point3(0), point3(165, 330, 165), white)));
    world.push_back(std::make_shared<translate>(vec3(50, 0, 100), box(point3(0), point3(165, 165, 165), white)));
    auto ql = Q(point3(343, 554, 332), vec3(-130, 0, 0), vec3(0, 0, -105), emit(15));
    world.push_back(ql);
    light = ql;
    use_bvh = true;
    bool shipped = name == "cornell_shipped";
    cam.initialize_perspective(pick(a.width, 600), 1.0, point3(278, 278, -800), point3(278, 278, 0), 1, 40.0, pick(a.spp, shipped ? 40 : 100),
                               pick(a.depth, shipped ? 4 : 50));
    cam.background_ = solid_color::black;
  } else if (name == "rtow") {
    // C2 (configs[1]): composed from src/main.cc:105-147 with the stationary sphere ctor (src/sphere.h:7), SURVEY.md §8d
    world.push_back(S(point3(0, -1000, 0), 1000, lamt(std::make_shared<checker_texture>(color(1, 1, 1), color(.6, .6, .2), 1.0))));
    for (int gx = -11; gx < 11; gx++)
      for (int gz = -11; gz < 11; gz++) {
        double choose = random_double();
        point3 c(gx + 0.9 * random_double(), 0.2, gz + 0.9 * random_double());
        if ((c - point3(4, 0.2, 0)).length() <= 0.9) continue;
        if (choose < 0.8) {
          color alb = random_vec() * random_vec();
          world.push_back(S(c, 0.2, std::make_shared<lambertian>(alb)));
        } else if (choose < 0.95) {
          color alb = random_vec(0.5, 1);
          double fuzz = random_double(0, 0.5);
          world.push_back(S(c, 0.2, std::make_shared<metal>(alb, fuzz)));
        } else {
          world.push_back(S(c, 0.2, std::make_shared<dielectric>(1.5)));
        }
      }
    world.push_back(S(point3(0, 1, 0), 1.0, std::make_shared<dielectric>(1.5)));
    world.push_back(S(point3(-4, 1, 0), 1.0, lam(.4, .2, .1)));
    world.push_back(S(point3(4, 1, 0), 1.0, std::make_shared<metal>(color(.7, .6, .5), 0.0)));
    use_bvh = true;
    cam.initialize_lens(pick(a.width, 1200), 16.0 / 9.0, point3(13, 2, 3), point3(0, 0, 0), 0.6, 10, 20, pick(a.spp, 500), pick(a.depth, 50));
    cam.background_ = solid(.7, .8, 1.0);
  } else if (name == "rtnw") {
    // C3 (configs[2]): composed from src/main.cc:155-171, 227-253, 402-437 per SURVEY.md §8d
    hittable_list boxes;
    auto ground = lam(0.48, 0.83, 0.53);
    for (int i = 0; i < 20; i++)
      for (int j = 0; j < 20; j++) {
        double w = 100.0, x0 = -1000.0 + i * w, z0 = -1000.0 + j * w;
        boxes.push_back(box(point3(x0, 0.0, z0), point3(x0 + w, random_double(1, 101), z0 + w), ground));
      }
    world.push_back(std::make_shared<bvh_node>(boxes));
    auto ql = Q(point3(123, 554, 147), vec3(300, 0, 0), vec3(0, 0, 265), emit(7));
    world.push_back(ql);
    light = ql;
    world.push_back(S(point3(400, 400, 200), 50, lam(0.7, 0.3, 0.1)));
    world.push_back(S(point3(260, 150, 45), 50, std::make_shared<dielectric>(1.5)));
    world.push_back(S(point3(0, 150, 145), 50, std::make_shared<metal>(color(0.8, 0.8, 0.9), 1.0)));
    auto glass_ball = S(point3(360, 150, 145), 70, std::make_shared<dielectric>(1.5));
    world.push_back(glass_ball);
    world.push_back(std::make_shared<volumne>(glass_ball, 0.2, solid(0.2, 0.4, 0.9)));
    world.push_back(std::make_shared<volumne>(S(point3(0, 0, 0), 5000, std::make_shared<dielectric>(1.5)), 0.0001, solid(1)));
    world.push_back(S(point3(400, 200, 400), 100, lamt(picture(A + "/earthmap.jpg"))));
    world.push_back(S(point3(220, 280, 300), 80, lamt(std::make_shared<perlin_texture>(5))));
    hittable_list balls;
    auto white = lam(.73, .73, .73);
    for (int j = 0; j < 1000; j++) balls.push_back(S(random_vec(0, 165), 10, white));
    world.push_back(std::make_shared<translate>(vec3(-100, 270, 395), std::make_shared<rotate_y>(std::make_shared<bvh_node>(balls), 15)));
    use_bvh = true;
    cam.initialize_perspective(pick(a.width, 800), 1.0, point3(478, 278, -600), point3(278, 278, 0), 1, 40.0, pick(a.spp, 1000), pick(a.depth, 40));
  } else if (name == "fox" || name == "glass_fox") {
    // C4 (configs[3]) = triangles of src/main.cc:346-393 under a fisheye camera; "glass_fox" = the shipped perspective view (:396)
    for (auto &t : load_gltf_triangles(A + "/Fox/glTF/Fox.gltf"))
      world.push_back(std::make_shared<triangle>(t[0], t[1], t[2], std::make_shared<dielectric>(solid(1.0), 1.5)));
    use_bvh = true;
    if (name == "fox")
      cam.initialize_fisheye(pick(a.width, 1920), 16.0 / 9.0, point3(220, 220, 220), point3(0, 20, 0), 1, 90, pick(a.spp, 1024), pick(a.depth, 5));
    else
      cam.initialize_perspective(pick(a.width, 600), 1.0, point3(220, 220, 220), point3(0, 20, 0), 1, 45.0, pick(a.spp, 200), pick(a.depth, 5));
    cam.background_ = picture(A + "/bathroom.exr");
  } else if (name == "sponza") {
    // C5 (configs[4]): src/main.cc:440-497 with caller-supplied triangles (Sponza.bin is missing: SURVEY.md §8c)
    if (!a.mesh || a.n_mesh_tris <= 0) throw std::runtime_error("scene 'sponza' needs a triangle array");
    for (int i = 0; i < a.n_mesh_tris; i++) {
      const float *p = a.mesh + 9 * (size_t)i;
      world.push_back(std::make_shared<triangle>(vec3(p[0], p[1], p[2]), vec3(p[3], p[4], p[5]), vec3(p[6], p[7], p[8]),
                                                 std::make_shared<lambertian>(solid(1.0))));
    }
    auto ql = Q(point3(0, 1200, 0), vec3(500, 0, 0), vec3(0, 0, 500), emit(10));
    world.push_back(ql);
    light = ql;
    use_bvh = true;
    cam.initialize_perspective(pick(a.width, 3840), 16.0 / 9.0, point3(500, 320, 90), point3(0, 280, 0), 1, 45.0, pick(a.spp, 4096), pick(a.depth, 5));
  } else if (name == "three_ball" || name == "three_ball_lens") {
    // src/main.cc:69-103
    world.push_back(S(point3(0, -1000, 0), 1000, lamt(std::make_shared<checker_texture>(color(1, 1, 1), color(.6, .6, .2), 1.0))));
    world.push_back(S(point3(0, 1, 0), 1.0, std::make_shared<dielectric>(solid(1), 1.5)));
    world.push_back(S(point3(-4, 1, 0), 1.0, lam(.4, .2, .1)));
    world.push_back(S(point3(4, 1, 0), 1.0, std::make_shared<metal>(solid(.7, .6, .5), 0.0)));
    if (name == "three_ball")
      cam.initialize_perspective(pick(a.width, 1280), 16.0 / 9.0, point3(13, 2, 3), vec3(0, 0, 0), 1, 20.0, pick(a.spp, 100), pick(a.depth, 5));
    else
      cam.initialize_lens(pick(a.width, 1280), 16.0 / 9.0, point3(13, 2, 3), vec3(1, 1, 1), 2.0, 15, 20.0, pick(a.spp, 1000), pick(a.depth, 5));
    cam.background_ = solid(.7, .8, 1.0);
  } else if (name == "motion_ball") {
    // src/main.cc:105-153 (moving spheres: reproduces the center_ quirk)
    world.push_back(S(point3(0, -1000, 0), 1000, lamt(std::make_shared<checker_texture>(color(1, 1, 1), color(.6, .6, .2), 1.0))));
    for (int gx = -11; gx < 11; gx++)
      for (int gz = -11; gz < 11; gz++) {
        double choose = random_double();
        point3 c1(gx + 0.7 * random_double(), 0.2, gz + 0.7 * random_double());
        point3 c2 = c1 + vec3(0, random_double(0, .15), 0);
        if ((c1 - point3(4, 0.2, 0)).length() <= 0.9 || choose < 0.3) continue;
        mptr m;
        if (choose < 0.8)
          m = lamt(std::make_shared<solid_color>(random_vec() * random_vec()));
        else if (choose < 0.95)
          m = std::make_shared<metal>(std::make_shared<solid_color>(random_vec(0.5, 1)), 0.0);
        else
          m = std::make_shared<dielectric>(solid(1), 1.5);
        world.push_back(std::make_shared<sphere>(c1, c2, 0.2, m));
      }
    auto glass = std::make_shared<dielectric>(solid(1), 1.5);
    world.push_back(S(point3(0, 1, 0), 1.0, glass));
    world.push_back(S(point3(-4, 1, 0), 1.0, lam(.4, .2, .1)));
    world.push_back(S(point3(4, 1, 0), 1.0, glass));
    use_bvh = true;
    cam.initialize_perspective(pick(a.width, 1280), 16.0 / 9.0, point3(13, 2, 3), point3(0, 0, 0), 1, 20, pick(a.spp, 20), pick(a.depth, 50));
    cam.background_ = solid(.7, .8, 1.0);
  } else if (name == "light_earth") {
    // src/main.cc:155-171
    world.push_back(S(point3(0, -1000, 0), 1000, lamt(std::make_shared<perlin_texture>(4))));
    world.push_back(S(point3(0, 2, 0), 2, std::make_shared<gloss>(picture(A + "/earthmap.jpg"), 1.0, 0.08)));
    auto ql = Q(point3(-2, 7, -2), vec3(4, 0, 0), vec3(0, 0, 4), emit(9));
    world.push_back(ql);
    light = ql;
    cam.initialize_perspective(pick(a.width, 1280), 16.0 / 9.0, point3(26, 3, 6), point3(0, 2, 0), 1, 20.0, pick(a.spp, 500), pick(a.depth, 5));
    cam.background_ = solid_color::black;
  } else if (name == "skybox_fisheye") {
    // src/main.cc:173-183
    world.push_back(S(vec3(0), 1, std::make_shared<dielectric>(solid_color::white, 1.0)));
    cam.initialize_fisheye(pick(a.width, 600), 1, point3(1.1, 1.8, 1.1), point3(0, 0, 0), 1.0, 90, pick(a.spp, 500), pick(a.depth, 5));
    cam.background_ = picture(A + "/bathroom.exr");
  } else if (name == "skybox_motion") {
    // src/main.cc:185-196
    world.push_back(std::make_shared<sphere>(vec3(-0.2, 0, 0), vec3(0.2, 0, 0), 1, lamt(picture(A + "/earthmap.jpg"))));
    cam.initialize_perspective(pick(a.width, 600), 1, point3(0, 0, 4), point3(0, 0, 0), 1.0, 70, pick(a.spp, 500), pick(a.depth, 5));
    cam.background_ = picture(A + "/bathroom.exr");
  } else if (name == "cornell_volume" || name == "cornell_specular" || name == "cornell_rotated") {
    // src/main.cc:227-307
    auto red = lam(.65, .05, .05), white = lam(.73, .73, .73), green = lam(.12, .45, .15);
    cornell_shell_b(world, red, white, green);
    auto yrot_box = [&](vec3 off, double sx, double sy, double sz, double deg, mptr m) {
      return std::make_shared<translate>(off, std::make_shared<rotate_y>(box(point3(0), point3(sx, sy, sz), m), deg));
    };
    int dflt_spp = 100;
    if (name == "cornell_volume") {
      world.push_back(std::make_shared<volumne>(yrot_box(vec3(265, 0, 285), 150, 280, 150, 45, white), 0.02, solid(0)));
      world.push_back(std::make_shared<volumne>(yrot_box(vec3(130, 0, 65), 140, 140, 140, -15, white), 0.02, solid(1)));
    } else if (name == "cornell_specular") {
      world.push_back(yrot_box(vec3(265, 0, 285), 150, 280, 150, 20, std::make_shared<metal>(color(1.0), 0.0)));
      world.push_back(yrot_box(vec3(130, 0, 65), 140, 140, 140, -15, white));
      dflt_spp = 500;
    } else {
      world.push_back(std::make_shared<rotate_z>(box(point3(265, 0, 295), point3(430, 330, 460), white), 15));
    }
    auto ql = Q(point3(113, 554, 127), vec3(330, 0, 0), vec3(0, 0, 305), emit(7));
    world.push_back(ql);
    light = ql;
    cam.initialize_perspective(pick(a.width, 600), 1.0, point3(278, 278, -800), point3(278, 278, 0), 1, 40, pick(a.spp, dflt_spp), pick(a.depth, 5));
    cam.background_ = solid_color::black;
  } else if (name == "perlin_ball") {
    // src/main.cc:402-437
    hittable_list boxes;
    auto ground = lam(0.48, 0.83, 0.53);
    for (int i = 0; i < 20; i++)
      for (int j = 0; j < 20; j++) {
        double w = 100.0, x0 = -1000.0 + i * w, z0 = -1000.0 + j * w;
        boxes.push_back(box(point3(x0, 0.0, z0), point3(x0 + w, random_double(1, 101), z0 + w), ground));
      }
    world.push_back(std::make_shared<bvh_node>(boxes));
    world.push_back(Q(point3(123, 554, 147), vec3(300, 0, 0), vec3(0, 0, 265), emit(7)));
    world.push_back(S(point3(260, 150, 45), 50, std::make_shared<dielectric>(1.5)));
    world.push_back(std::make_shared<translate>(
        point3(180, 280, 400), std::make_shared<rotate_x>(S(point3(0), 80, lamt(std::make_shared<perlin_texture>(8))), -90)));
    use_bvh = true;
    cam.initialize_perspective(pick(a.width, 600), 1.0, point3(478, 278, -600), point3(278, 278, 0), 1, 40.0, pick(a.spp, 500), pick(a.depth, 5));
  } else if (name == "sphereflake") {
    // src/main.cc:48-67 (renders the plain list: the BVH is built but unused, :62-63)
    sphereflake_gen(100, point3(0, 0, 0), std::make_shared<metal>(solid(.5)), 4, world, 0.25, vec3(0, 1, 0));
    cam.initialize_perspective(pick(a.width, 400), 1.0, point3(200), vec3(0, 0, 0), 1, 90.0, pick(a.spp, 50), pick(a.depth, 5));
    cam.background_ = picture(A + "/bathroom.exr");
  } else if (name.rfind("zoo_", 0) == 0) {
    // Not a scene of src/main.cc: every material (src/material.h:57-219) side by side under each kind of `light` argument that
    // camera::render accepts (src/camera.h:135): none, a quad (src/quad.h:66-78), a sphere (src/sphere.h:76-81).
    world.push_back(S(point3(0, -1000, 0), 1000, lamt(std::make_shared<checker_texture>(color(.9, .9, .9), color(.2, .3, .1), 0.8))));
    world.push_back(S(point3(-4, 1, 0), 1.0, lam(.4, .2, .1)));
    world.push_back(S(point3(-2, 1, 0), 1.0, std::make_shared<metal>(color(.7, .6, .5), 0.3)));
    world.push_back(S(point3(0, 1, 0), 1.0, std::make_shared<dielectric>(1.5)));
    world.push_back(S(point3(2, 1, 0), 1.0, std::make_shared<gloss>(color(.8, .3, .3), 0.6, 0.5)));
    world.push_back(std::make_shared<volumne>(S(point3(4, 1, 0), 1.0, lam(1, 1, 1)), 0.8, solid(0.2, 0.4, 0.9)));
    world.push_back(std::make_shared<translate>(vec3(-1.5, 0, 2.2), std::make_shared<rotate_y>(box(point3(0), point3(1, 1.4, 1), lam(.73, .73, .73)), 25)));
    world.push_back(S(point3(1.2, 0.5, 2.6), 0.5, std::make_shared<gloss>(std::make_shared<perlin_texture>(3), 0.2, 0.1)));
    auto ql = Q(point3(-1, 4, -1), vec3(2, 0, 0), vec3(0, 0, 2), emit(8));
    auto sl = std::make_shared<sphere>(point3(3, 3.5, -2), 0.6, emit(10));
    world.push_back(ql);
    world.push_back(sl);
    if (name == "zoo_quad")
      light = ql;
    else if (name == "zoo_sphere")
      light = sl;
    else if (name != "zoo_nolight")
      throw std::runtime_error("unknown scene " + name);
    use_bvh = true;
    cam.initialize_perspective(pick(a.width, 400), 16.0 / 9.0, point3(0, 3, 9), point3(0, 1, 0), 1, 40.0, pick(a.spp, 100), pick(a.depth, 10));
    cam.background_ = solid(.10, .12, .20);
  } else if (name == "white_sphere") {
    // src/main.cc:500-510
    world.push_back(S(point3(0, 0, 0), 1, std::make_shared<metal>(solid(1.0), 0.1)));
    cam.initialize_perspective(pick(a.width, 400), 1.0, point3(13, 2, 3), point3(0, 0, 0), 1, 20, pick(a.spp, 100), pick(a.depth, 5));
    cam.background_ = solid(1.0);
  } else if (name == "fuzz_metal" || name == "glossy_ball") {
    // src/main.cc:512-547 and 309-343
    gallery_shell(world);
    if (name == "fuzz_metal") {
      const double xs[5] = {2, 5.5, 9, 12.5, 16};
      const float fz[5] = {0.0f, 0.25f, 0.5f, 0.75f, 1.0f};
      for (int i = 0; i < 5; i++) world.push_back(S(point3(xs[i], 0, -0.5), 1.25, std::make_shared<metal>(solid_color::white, fz[i])));
    } else {
      auto earth = picture(A + "/earthmap.jpg");
      const double xs[4] = {3, 7, 11, 15};
      const float pr[4] = {1.00f, 0.40f, 0.15f, 0.02f};
      for (int i = 0; i < 4; i++) world.push_back(S(point3(xs[i], 0, -0.5), 1.25, std::make_shared<gloss>(earth, 1.0, pr[i])));
    }
    auto ql = Q(point3(5.5, 3.995, -1.25), vec3(7, 0, 0), vec3(0, 0, 2.5), emit(name == "fuzz_metal" ? 7 : 8));
    world.push_back(ql);
    light = ql;
    use_bvh = true;
    cam.initialize_perspective(pick(a.width, 760), 19.0 / 9.0, point3(9, 0, 15.2), point3(9, 0, 1), 1, 40.0, pick(a.spp, 1000), pick(a.depth, 10));
    cam.background_ = solid_color::black;
  } else if (name == "infinite_reflection") {
    // src/main.cc:549-579
    auto red = lam(.65, .05, .05), white = lam(.73, .73, .73), green = lam(.12, .45, .15);
    auto mirror = std::make_shared<metal>(solid(0.80), 0.0);
    world.push_back(Q(point3(555, 0, 0), vec3(0, 555, 0), vec3(0, 0, 555), green));
    world.push_back(Q(point3(0, 0, 0), vec3(0, 555, 0), vec3(0, 0, 555), red));
    world.push_back(Q(point3(0, 0, 0), vec3(555, 0, 0), vec3(0, 0, 555), white));
    world.push_back(Q(point3(555, 555, 555), vec3(-555, 0, 0), vec3(0, 0, -555), white));
    world.push_back(Q(point3(0, 0, 555), vec3(555, 0, 0), vec3(0, 555, 0), mirror));
    world.push_back(Q(point3(0, 0, 0), vec3(555, 0, 0), vec3(0, 555, 0), mirror));
    world.push_back(S(point3(460, 80, 80), 60, std::make_shared<gloss>(picture(A + "/earthmap.jpg"), 0.97, 0.18)));
    world.push_back(std::make_shared<translate>(vec3(130, 0, 65), std::make_shared<rotate_y>(box(point3(0), point3(140, 140, 140), white), -15)));
    auto ql = Q(point3(113, 554, 127), vec3(330, 0, 0), vec3(0, 0, 305), emit(5));
    world.push_back(ql);
    light = ql;
    use_bvh = true;
    cam.initialize_perspective(pick(a.width, 600), 1.0, point3(500, 290, 550), point3(400, 278, 0), 1, 40.0, pick(a.spp, 1000), pick(a.depth, 30));
    cam.background_ = solid_color::black;
  } else if (name.rfind("noise_", 0) == 0) {
    // src/main.cc:581-631: one textured quad under the orthographic camera
    tptr t;
    double side = 40, vh = 20;
    if (name == "noise_perlin") {
      t = std::make_shared<perlin_texture>(1);
      side = 10;
      vh = 10;
    } else if (name == "noise_value")
      t = std::make_shared<value_texture>(40);
    else if (name == "noise_worley")
      t = std::make_shared<worley_texture>();
    else if (name == "noise_voronoi")
      t = std::make_shared<voronoi_texture>();
    else
      throw std::runtime_error("unknown scene " + name);
    world.push_back(Q(point3(0, 0, 0), vec3(side, 0, 0), vec3(0, side, 0), lamt(t)));
    cam.initialize_orthnormal(pick(a.width, 400), 1, vh, vec3(side / 2, side / 2, 1), vec3(side / 2, side / 2, 0), pick(a.spp, 10), pick(a.depth, 5));
    cam.background_ = solid_color::white;
  } else {
    throw std::runtime_error("unknown scene " + name);
  }

  if (use_bvh)
    sc.world = std::make_shared<bvh_node>(world);
  else
    sc.world = std::make_shared<hittable_list>(world);
  sc.light = light;
}

std::mutex g_mu;
std::map<int, std::unique_ptr<scene>> g_scenes;
int g_next = 1;
thread_local std::string t_err;

scene *get(int h) {
  std::lock_guard<std::mutex> lk(g_mu);
  auto it = g_scenes.find(h);
  return it == g_scenes.end() ? nullptr : it->second.get();
}

struct cout_silencer { // camera::render prints one line per row (src/camera.h:161)
  std::streambuf *old;
  std::ostringstream sink;
  cout_silencer() : old(std::cout.rdbuf(sink.rdbuf())) {}
  ~cout_silencer() { std::cout.rdbuf(old); }
};

} // namespace

extern "C" {

const char *ref_last_error() { return t_err.c_str(); }

// Build a scene through the reference API.  construct_seed -> srand() before construction (scene
// construction consumes std::rand: src/main.cc:112-127,413; src/noise.h:14,87).  Returns a handle > 0, or -1.
int ref_scene_create(const char *name, int width, int spp, int depth, const char *asset_dir, unsigned construct_seed, const float *mesh,
                     int n_mesh_tris) {
  try {
    auto sc = std::make_unique<scene>();
    sc->name = name;
    g_rng_mode = 0;
    srand(construct_seed);
    build_args a{width, spp, depth, asset_dir ? asset_dir : ".", mesh, n_mesh_tris};
    {
      cout_silencer quiet;
      build(*sc, name, a);
    }
    // tag leaves (ids in DFS order), then flatten the live objects
    instrumenter ins;
    hptr root = sc->world;
    ins.descend(root.get());
    sc->flat.run(*sc->world, sc->light.get(), sc->cam.background_.get());
    sc->desc = sc->flat.out.desc();
    sc->cam_desc = b200::flatten_camera(sc->cam);
    sc->tex_by_row.assign(sc->flat.out.textures.size(), nullptr);
    for (auto &kv : sc->flat.tex_rows_)
      if (!sc->tex_by_row[kv.second]) sc->tex_by_row[kv.second] = static_cast<const texture *>(kv.first);
    std::lock_guard<std::mutex> lk(g_mu);
    int h = g_next++;
    g_scenes[h] = std::move(sc);
    return h;
  } catch (const std::exception &e) {
    t_err = e.what();
    return -1;
  }
}

// Flatten the same live objects again with the adaptor's texture opt-ins (include/b200rt_adaptor.hpp: flattener::float_images,
// ::bilinear_pictures); the description returned by ref_scene_desc changes accordingly.  0/0 restores the reference semantics.
int ref_scene_reflatten(int h, int float_images, int bilinear) {
  scene *s = get(h);
  if (!s) return -1;
  try {
    s->flat = b200::flattener();
    s->flat.float_images = float_images != 0;
    s->flat.bilinear_pictures = bilinear != 0;
    s->flat.run(*s->world, s->light.get(), s->cam.background_.get());
    s->desc = s->flat.out.desc();
    return 0;
  } catch (const std::exception &e) {
    t_err = e.what();
    return -1;
  }
}

void ref_scene_destroy(int h) {
  std::lock_guard<std::mutex> lk(g_mu);
  g_scenes.erase(h);
}

const b200rt_scene_desc *ref_scene_desc(int h) {
  scene *s = get(h);
  return s ? &s->desc : nullptr;
}

int ref_scene_camera(int h, b200rt_camera *out) {
  scene *s = get(h);
  if (!s) return -1;
  *out = s->cam_desc;
  return 0;
}

// camera::generate_ray (src/camera.h:244-284) for `spp` samples of every pixel, after srand(seed);
// order: index = k * W*H + (i*W + j).
int ref_generate_rays(int h, unsigned seed, int spp, double *rays7) {
  scene *s = get(h);
  if (!s) return -1;
  camera &cam = s->cam;
  g_rng_mode = 0;
  srand(seed);
  point3 du = cam.viewport_width_ * cam.right_ / cam.image_width_;    // src/camera.h:137
  point3 dv = -cam.viewport_height_ * cam.up_ / cam.image_height_;    // src/camera.h:138
  size_t npix = (size_t)cam.image_width_ * cam.image_height_;
  for (int k = 0; k < spp; k++)
    for (int i = 0; i < cam.image_height_; i++)
      for (int j = 0; j < cam.image_width_; j++) {
        ray r = cam.generate_ray(i, j, du, dv);
        double *o = rays7 + 7 * (k * npix + (size_t)i * cam.image_width_ + j);
        for (int c = 0; c < 3; c++) o[c] = r.origin().e[c], o[3 + c] = r.direction().e[c];
        o[6] = r.time();
      }
  return 0;
}

// Same, also returning the std::rand values each generate_ray call consumed, in consumption order
// (as random_double values, src/utility.h:20): draws[i*max_draws + k], counts[i].  Lets the restatement
// be checked bit for bit on identical draws.
int ref_generate_rays_logged(int h, unsigned seed, int spp, double *rays7, double *draws, int max_draws, int32_t *counts) {
  scene *s = get(h);
  if (!s) return -1;
  camera &cam = s->cam;
  g_rng_mode = 0;
  srand(seed);
  point3 du = cam.viewport_width_ * cam.right_ / cam.image_width_;
  point3 dv = -cam.viewport_height_ * cam.up_ / cam.image_height_;
  size_t npix = (size_t)cam.image_width_ * cam.image_height_;
  std::vector<int> log;
  t_rand_log = &log;
  for (int k = 0; k < spp; k++)
    for (int i = 0; i < cam.image_height_; i++)
      for (int j = 0; j < cam.image_width_; j++) {
        log.clear();
        ray r = cam.generate_ray(i, j, du, dv);
        size_t idx = k * npix + (size_t)i * cam.image_width_ + j;
        double *o = rays7 + 7 * idx;
        for (int c = 0; c < 3; c++) o[c] = r.origin().e[c], o[3 + c] = r.direction().e[c];
        o[6] = r.time();
        counts[idx] = (int32_t)log.size();
        for (int q = 0; q < max_draws; q++) draws[idx * max_draws + q] = q < (int)log.size() ? log[q] / (RAND_MAX + 1.0) : 0.0;
      }
  t_rand_log = nullptr;
  return 0;
}

// <world>.hit(r, interval(tmin, tmax), record) on a batch (src/camera.h:198).  Volumes draw from std::rand after srand(seed).
int ref_trace(int h, const double *rays7, size_t n, double tmin, double tmax, unsigned seed, int32_t *prim, double *t, double *uv,
              uint8_t *front, double *p, double *normal) {
  scene *s = get(h);
  if (!s) return -1;
  g_rng_mode = 0;
  srand(seed);
  auto work = [&](size_t lo, size_t hi) {
    for (size_t i = lo; i < hi; i++) {
      const double *q = rays7 + 7 * i;
      ray r(point3(q[0], q[1], q[2]), vec3(q[3], q[4], q[5]), q[6]);
      hit_record rec;
      rec.t = 0;
      rec.u = rec.v = 0; // triangle::hit and volumne::hit leave u,v unwritten (src/triangle.h:35-39); the ABI defines them as 0
      rec.front_face = false;
      b200::leaf_tag::last_hit_id() = -1;
      bool ok = s->world->hit(r, interval(tmin, tmax), rec);
      if (prim) prim[i] = ok ? b200::leaf_tag::last_hit_id() : -1;
      if (t) t[i] = ok ? rec.t : 0.0;
      if (uv) uv[2 * i] = ok ? rec.u : 0.0, uv[2 * i + 1] = ok ? rec.v : 0.0;
      if (front) front[i] = ok ? (rec.front_face ? 1 : 0) : 0;
      for (int c = 0; c < 3; c++) {
        if (p) p[3 * i + c] = ok ? rec.p.e[c] : 0.0;
        if (normal) normal[3 * i + c] = ok ? rec.normal.e[c] : 0.0;
      }
    }
  };
  // hit() is const and the last-hit id is thread-local: without volumes (the only std::rand consumer under hit, src/volumne.h:36)
  // rays are independent and a large batch is spread over the host threads
  const unsigned hw = std::max(1u, std::thread::hardware_concurrency());
  if (n < 65536 || s->desc.n_volumes > 0 || hw == 1) {
    work(0, n);
  } else {
    std::vector<std::thread> pool;
    for (unsigned k = 0; k < hw; k++) pool.emplace_back(work, n * k / hw, n * (k + 1) / hw);
    for (auto &th : pool) th.join();
  }
  return 0;
}

// texture::sample(u, v, p) of texture row `tex` (rows as in the flattened description).
int ref_texture_sample(int h, int tex, const double *uvp5, size_t n, double *rgb) {
  scene *s = get(h);
  if (!s || tex < 0 || tex >= (int)s->tex_by_row.size() || !s->tex_by_row[tex]) return -1;
  texture *t = const_cast<texture *>(s->tex_by_row[tex]);
  for (size_t i = 0; i < n; i++) {
    const double *q = uvp5 + 5 * i;
    color c = t->sample(q[0], q[1], point3(q[2], q[3], q[4]));
    for (int k = 0; k < 3; k++) rgb[3 * i + k] = c.e[k];
  }
  return 0;
}

// camera::ray_color(r, world, depth, light) on a batch, after srand(seed) (src/camera.h:193-241).
int ref_ray_color(int h, const double *rays7, size_t n, int depth, unsigned seed, double *rgb) {
  scene *s = get(h);
  if (!s) return -1;
  g_rng_mode = 0;
  srand(seed);
  for (size_t i = 0; i < n; i++) {
    const double *q = rays7 + 7 * i;
    ray r(point3(q[0], q[1], q[2]), vec3(q[3], q[4], q[5]), q[6]);
    color c = s->cam.ray_color(r, *s->world, depth, s->light);
    for (int k = 0; k < 3; k++) rgb[3 * i + k] = c.e[k];
  }
  return 0;
}

// One level of camera::ray_color below the hit test, through the reference's OWN objects, on a batch of rays, with every
// std::rand value the level consumes logged (src/camera.h:202-240).  Per ray: world.hit (records a miss as mode -1; the
// free-flight draws of volumes are taken but not logged), then exactly what ray_color does: mat->emitted, mat->scatter, and
// for scatter_mode::kRandom the pdf it builds (scatter_record.pdf, or dual_pdf(hittable_pdf(light), scatter_record.pdf)),
// generate(), value(), mat->p_scattered.
//   rec12  : {p[3], normal[3], t, u, v, front_face, material row of the flattened description, primitive id}
//   out16  : {mode, emitted[3], attenuation[3], direction[3], pdf_value, p_scattered, weight[3], 0}  (layout of b200rt_scatter)
//   draws  : n x max_draws random_double values in consumption order; counts[i] = how many
int ref_scatter_logged(int h, const double *rays7, size_t n, unsigned seed, double *rec12, double *out16, double *draws, int max_draws, int32_t *counts) {
  scene *s = get(h);
  if (!s) return -1;
  g_rng_mode = 0;
  srand(seed);
  std::vector<int> log;
  for (size_t i = 0; i < n; i++) {
    const double *q = rays7 + 7 * i;
    ray r(point3(q[0], q[1], q[2]), vec3(q[3], q[4], q[5]), q[6]);
    double *hr = rec12 + 12 * i, *o = out16 + 16 * i;
    for (int k = 0; k < 12; k++) hr[k] = 0;
    for (int k = 0; k < 16; k++) o[k] = 0;
    for (int k = 0; k < max_draws; k++) draws[i * max_draws + k] = 0;
    counts[i] = 0;
    hit_record record;
    record.t = 0, record.u = record.v = 0, record.front_face = false;
    b200::leaf_tag::last_hit_id() = -1;
    if (!s->world->hit(r, interval(0.001, infinity), record)) {
      o[0] = -1;
      continue;
    }
    auto row = s->flat.mat_rows_.find(record.mat.get());
    if (row == s->flat.mat_rows_.end()) {
      t_err = "ref_scatter_logged: hit a material the flattener has not seen";
      return -2;
    }
    for (int c = 0; c < 3; c++) hr[c] = record.p.e[c], hr[3 + c] = record.normal.e[c];
    hr[6] = record.t, hr[7] = record.u, hr[8] = record.v, hr[9] = record.front_face ? 1.0 : 0.0;
    hr[10] = row->second, hr[11] = b200::leaf_tag::last_hit_id();
    log.clear();
    t_rand_log = &log;
    color emitted = record.mat->emitted(r, record, record.u, record.v, record.p); // src/camera.h:205
    scatter_record srec;
    color weight(0);
    if (!record.mat->scatter(r, record, srec)) { // :209
      o[0] = 0;
    } else if (srec.mode == scatter_mode::kDetermined) { // :212-216
      o[0] = 1;
      for (int c = 0; c < 3; c++) o[4 + c] = srec.attenuation.e[c], o[7 + c] = srec.ray_scattered.direction().e[c];
      weight = srec.attenuation;
    } else {
      o[0] = 2;
      vec3 dir;
      double pdf_value;
      if (s->light == nullptr) { // :219-222
        auto pdf = srec.pdf;
        dir = pdf->generate();
        pdf_value = pdf->value(dir);
      } else { // :229-234
        auto p0 = std::make_shared<hittable_pdf>(*s->light, record.p);
        auto p1 = srec.pdf;
        dual_pdf dual(p0, p1);
        dir = dual.generate();
        pdf_value = dual.value(dir);
      }
      ray scattered_ray(record.p, dir, r.time());
      double p_scattered = record.mat->p_scattered(r, record, scattered_ray); // :225, :237
      for (int c = 0; c < 3; c++) o[4 + c] = srec.attenuation.e[c], o[7 + c] = dir.e[c];
      o[10] = pdf_value, o[11] = p_scattered;
      weight = (srec.attenuation * p_scattered) / pdf_value; // the factor on input_color in :227 / :239, input_color taken as 1
    }
    t_rand_log = nullptr;
    for (int c = 0; c < 3; c++) o[1 + c] = emitted.e[c], o[12 + c] = weight.e[c];
    counts[i] = (int32_t)log.size();
    for (int k = 0; k < max_draws && k < (int)log.size(); k++) draws[i * max_draws + k] = log[k] / (RAND_MAX + 1.0);
  }
  t_rand_log = nullptr;
  return 0;
}

struct ref_render_stats {
  uint64_t samples;
  uint64_t rays;
  uint64_t nan_samples;
  double seconds; // wall clock around the sample loops only
  int threads;
  int rng_mode;
};

// The loop nest of camera::render (src/camera.h:163-170) over all rows, `spp` samples per pixel, calling the
// reference's generate_ray and ray_color.  sum / sumsq (W*H*3, may be NULL) receive per-pixel sums of the sample
// radiance and of its square (samples with a non-finite channel are skipped there and counted in nan_samples).
//   rng_mode 0: glibc std::rand, srand(seed), rows in order on ONE thread (bit-reproduces the stock executable's stream)
//   rng_mode 1: thread-local stream seeded per row from (seed,row); rows handed to `threads` std::threads.
int ref_render(int h, unsigned seed, int spp, int threads, int rng_mode, double *sum, double *sumsq, ref_render_stats *st) {
  scene *s = get(h);
  if (!s) return -1;
  camera &cam = s->cam;
  const int W = cam.image_width_, H = cam.image_height_;
  if (rng_mode == 0) threads = 1;
  if (threads < 1) threads = 1;
  g_rng_mode = rng_mode;
  if (rng_mode == 0) srand(seed);
  point3 du = cam.viewport_width_ * cam.right_ / cam.image_width_;
  point3 dv = -cam.viewport_height_ * cam.up_ / cam.image_height_;
  counting_world cw(*s->world);
  std::atomic<int> next_row{0};
  std::atomic<uint64_t> rays{0}, nans{0};
  auto worker = [&]() {
    counting_world::counter() = 0;
    uint64_t my_nans = 0;
    for (;;) {
      int i = next_row.fetch_add(1);
      if (i >= H) break;
      if (rng_mode == 1) tls_seed(seed, (uint64_t)i);
      for (int j = 0; j < W; j++) {
        double a[3] = {0, 0, 0}, b[3] = {0, 0, 0};
        for (int k = 0; k < spp; k++) {
          ray r = cam.generate_ray(i, j, du, dv);
          color c = cam.ray_color(r, cw, cam.max_recur_depth_, s->light);
          if (!(std::isfinite(c.e[0]) && std::isfinite(c.e[1]) && std::isfinite(c.e[2]))) {
            my_nans++;
            continue;
          }
          for (int q = 0; q < 3; q++) a[q] += c.e[q], b[q] += c.e[q] * c.e[q];
        }
        size_t o = 3 * ((size_t)i * W + j);
        for (int q = 0; q < 3; q++) {
          if (sum) sum[o + q] = a[q];
          if (sumsq) sumsq[o + q] = b[q];
        }
      }
    }
    rays += counting_world::counter();
    nans += my_nans;
  };
  auto t0 = std::chrono::steady_clock::now();
  if (threads == 1) {
    worker();
  } else {
    std::vector<std::thread> pool;
    for (int t = 0; t < threads; t++) pool.emplace_back(worker);
    for (auto &t : pool) t.join();
  }
  auto t1 = std::chrono::steady_clock::now();
  g_rng_mode = 0;
  if (st) {
    st->samples = (uint64_t)W * H * spp;
    st->rays = rays.load();
    st->nan_samples = nans.load();
    st->seconds = std::chrono::duration<double>(t1 - t0).count();
    st->threads = threads;
    st->rng_mode = rng_mode;
  }
  return 0;
}

// camera::render exactly as written (src/camera.h:135-176): PSTL row loop (serial here: no TBB), P3 text to `ppm_path`
// (NULL -> /dev/null).  Returns wall seconds of the whole call; rgb_linear (may be NULL) receives cam.image_.
double ref_render_stock(int h, unsigned seed, int spp, const char *ppm_path, double *rgb_linear) {
  scene *s = get(h);
  if (!s) return -1.0;
  camera &cam = s->cam;
  int keep = cam.samples_per_pixel_;
  if (spp > 0) cam.samples_per_pixel_ = spp;
  g_rng_mode = 0;
  srand(seed);
  std::ofstream out(ppm_path ? ppm_path : "/dev/null");
  double secs;
  {
    cout_silencer quiet;
    auto t0 = std::chrono::steady_clock::now();
    cam.render(out, *s->world, s->light);
    secs = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
  }
  cam.samples_per_pixel_ = keep;
  if (rgb_linear)
    for (size_t i = 0; i < cam.image_.size(); i++)
      for (int k = 0; k < 3; k++) rgb_linear[3 * i + k] = cam.image_[i].e[k];
  return secs;
}

// The output stage of camera::render on caller-supplied pixel colours: the header lines of src/camera.h:149-151 and the
// reference's own write_color (src/color.h:22-36) once per pixel, into memory.  Returns the length (the text is truncated to cap).
size_t ref_write_colors(const double *rgb, int W, int H, char *out, size_t cap) {
  std::ostringstream os;
  int max = 255;
  os << "P3\n";
  os << W << ' ' << H << '\n';
  os << max << '\n';
  for (size_t i = 0; i < (size_t)W * H; i++) write_color(os, color(rgb[3 * i], rgb[3 * i + 1], rgb[3 * i + 2]));
  const std::string t = os.str();
  if (out && cap) std::memcpy(out, t.data(), std::min(cap, t.size()));
  return t.size();
}

// b200::render(cam, out, world, light) — the binding of INTEGRATION.md — in place of cam.render(out, world, light).
// `product_lib` = path of libb200rt.so, loaded here on first use.  Returns wall seconds, < 0 on error (ref_last_error()).
// device >= 0: that one device; device < 0: the first -device devices of the box (the frame split across them).
double ref_render_b200(int h, unsigned long long seed, int spp, int device, const char *ppm_path, double *rgb_linear, const char *product_lib) {
  scene *s = get(h);
  if (!s) return -1.0;
  if (!product_lib || !b200dyn::bind(product_lib, t_err)) return -2.0;
  camera &cam = s->cam;
  int keep = cam.samples_per_pixel_;
  if (spp > 0) cam.samples_per_pixel_ = spp;
  std::ofstream out(ppm_path ? ppm_path : "/dev/null");
  double secs = -3.0;
  try {
    auto t0 = std::chrono::steady_clock::now();
    std::vector<int> devices;
    if (device >= 0)
      devices.push_back(device);
    else
      for (int k = 0; k < -device; k++) devices.push_back(k);
    b200::render(cam, out, *s->world, s->light, devices, seed);
    secs = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    if (rgb_linear)
      for (size_t i = 0; i < cam.image_.size(); i++)
        for (int k = 0; k < 3; k++) rgb_linear[3 * i + k] = cam.image_[i].e[k];
  } catch (const std::exception &e) {
    t_err = e.what();
  }
  cam.samples_per_pixel_ = keep;
  return secs;
}

int ref_pstl_is_parallel() {
#if defined(_GLIBCXX_USE_TBB_PAR_BACKEND) && _GLIBCXX_USE_TBB_PAR_BACKEND
  return 1;
#else
  return 0;
#endif
}

} // extern "C"
