// b200rt_adaptor.hpp — reference-side C++ adaptor for libb200rt.so.
//
// This is the binding a maintainer of JTtNinjaCode/CPU-Ray-Tracing-Implementation adds to
// switch `cam.render(of, world, light)` (reference src/camera.h:135) over to the B200
// backend: it walks the reference's polymorphic scene graph (hittable / material /
// texture objects), emits the plain-old-data description of include/b200rt.h, and calls
// the C ABI.  It is compiled on the reference side, in the one translation unit that
// already includes the reference headers (they are not independently includable,
// SURVEY.md §1), so it contains no copy of reference code — only reads of its objects.
//
// Access to data members: the reference classes keep their state private and offer no
// getters (e.g. src/sphere.h:97-106).  Either add `friend class b200::flattener;` to
// those classes, or — for the untouched headers — include them like this:
//
//     #include <std headers the reference uses>      // before the macro, or libstdc++ breaks
//     #define private public
//     #include "aabb.h" ... "volumne.h"               // main.cc order, src/main.cc:5-20
//     #undef private
//     #include "b200rt_adaptor.hpp"
//
// Unknown hittable / material / texture subclasses raise std::runtime_error: there is no
// silent fallback to the CPU path.
#pragma once

#include <cstring>
#include <map>
#include <stdexcept>
#include <string>
#include <tuple>
#include <unordered_map>
#include <vector>

#include "b200rt.h"

namespace b200 {

// Instrumentation leaf used by the parity harness: forwards to the wrapped primitive and
// remembers which primitive produced the record that the reference returns.  Because every
// successful leaf hit overwrites the caller's record (src/hittable_list.h:24-28,
// src/bvh_node.h:51-54), the last successful leaf is the one whose record is returned.
class leaf_tag : public hittable {
public:
  leaf_tag(std::shared_ptr<hittable> inner, int id) : inner_(std::move(inner)), id_(id) {}
  bool hit(const ray &r, interval ray_t, hit_record &record) const override {
    if (!inner_->hit(r, ray_t, record)) return false;
    last_hit_id() = id_;
    return true;
  }
  aabb get_bounding_box() const override { return inner_->get_bounding_box(); }
  double pdf_value(const point3 &o, const vec3 &d) const override { return inner_->pdf_value(o, d); }
  vec3 random(const point3 &o) const override { return inner_->random(o); }
  static int &last_hit_id() {
    static thread_local int id = -1;
    return id;
  }
  std::shared_ptr<hittable> inner_;
  int id_;
};

struct flat_scene {
  std::vector<b200rt_node> nodes;
  std::vector<uint32_t> list_children;
  std::vector<double> spheres, quads, tris;
  std::vector<int32_t> sphere_mat, quad_mat, tri_mat;
  std::vector<b200rt_volume> volumes;
  std::vector<b200rt_material> materials;
  std::vector<b200rt_texture> textures;
  std::vector<b200rt_image> images;
  std::vector<uint8_t> image_bytes;
  std::vector<double> perlin_vec;
  std::vector<int32_t> perlin_perm;
  std::vector<b200rt_value_noise> value_noise;
  std::vector<float> value_lattice;
  uint32_t root = 0;
  uint32_t n_primitives = 0;
  int32_t light_kind = B200RT_LIGHT_NONE;
  int32_t light_index = -1;
  int32_t background_tex = -1;

  b200rt_scene_desc desc() const {
    b200rt_scene_desc d;
    std::memset(&d, 0, sizeof d);
    d.abi_version = B200RT_ABI_VERSION;
    d.root = root;
    d.n_nodes = (uint32_t)nodes.size();
    d.n_list_children = (uint32_t)list_children.size();
    d.nodes = nodes.data();
    d.list_children = list_children.data();
    d.n_spheres = (uint32_t)sphere_mat.size();
    d.n_quads = (uint32_t)quad_mat.size();
    d.n_tris = (uint32_t)tri_mat.size();
    d.n_volumes = (uint32_t)volumes.size();
    d.spheres = spheres.data();
    d.sphere_mat = sphere_mat.data();
    d.quads = quads.data();
    d.quad_mat = quad_mat.data();
    d.tris = tris.data();
    d.tri_mat = tri_mat.data();
    d.volumes = volumes.data();
    d.n_materials = (uint32_t)materials.size();
    d.n_textures = (uint32_t)textures.size();
    d.materials = materials.data();
    d.textures = textures.data();
    d.n_images = (uint32_t)images.size();
    d.n_perlin = (uint32_t)(perlin_perm.size() / 256);
    d.n_value_noise = (uint32_t)value_noise.size();
    d.n_primitives = n_primitives;
    d.images = images.data();
    d.image_bytes = image_bytes.data();
    d.image_bytes_size = image_bytes.size();
    d.perlin_vec = perlin_vec.data();
    d.perlin_perm = perlin_perm.data();
    d.value_noise = value_noise.data();
    d.value_lattice = value_lattice.data();
    d.value_lattice_size = value_lattice.size();
    d.light_kind = light_kind;
    d.light_index = light_index;
    d.background_tex = background_tex;
    return d;
  }
};

class flattener {
public:
  // Opt-ins that deliberately leave the reference's texture semantics (both default off = the reference's own lookups):
  bool float_images = false;      // hand image::fdata (un-quantised binary32) to the backend instead of the 8-bit image::bdata
  bool bilinear_pictures = false; // picture_texture lookups filter bilinearly instead of picking the nearest texel (src/texture.h:67-74)
public:
  flat_scene out;

  // world / light / background exactly as handed to camera::render (src/camera.h:135, :329).
  void run(const hittable &world, const hittable *light, const texture *background) {
    out = flat_scene();
    next_prim_ = 0;
    out.root = (uint32_t)visit(&world, false);
    out.n_primitives = (uint32_t)next_prim_;
    if (light) {
      const hittable *l = light;
      if (auto t = dynamic_cast<const leaf_tag *>(l)) l = t->inner_.get();
      if (auto q = dynamic_cast<const quad *>(l)) {
        out.light_kind = B200RT_LIGHT_QUAD;
        out.light_index = quad_row(q);
      } else if (auto s = dynamic_cast<const sphere *>(l)) {
        out.light_kind = B200RT_LIGHT_SPHERE;
        out.light_index = sphere_row(s);
      } else {
        out.light_kind = B200RT_LIGHT_DEFAULT; // hittable base pdf_value/random, src/hittable.h:39-41
      }
    }
    out.background_tex = background ? texture_row(background) : -1;
  }

  // bookkeeping (public so that test harnesses can map rows back to live objects)
  int next_prim_ = 0;
  std::unordered_map<const void *, int> sphere_rows_, quad_rows_, tri_rows_, prim_ids_, tex_rows_, mat_rows_, image_rows_;
  std::map<std::tuple<double, double, double>, int> solid_by_value_;
  std::map<std::tuple<int, int, float, float>, int> mat_by_value_;

  static void put3(std::vector<double> &v, const vec3 &a) {
    v.push_back(a.e[0]);
    v.push_back(a.e[1]);
    v.push_back(a.e[2]);
  }
  int new_node(int kind) {
    b200rt_node n;
    std::memset(&n, 0, sizeof n);
    n.kind = kind;
    n.a = n.b = 0;
    n.c = -1;
    out.nodes.push_back(n);
    return (int)out.nodes.size() - 1;
  }
  int prim_id(const void *obj, int forced, bool in_boundary) {
    if (in_boundary) return -1;
    auto it = prim_ids_.find(obj);
    if (it != prim_ids_.end()) return it->second;
    int id = forced >= 0 ? forced : next_prim_;
    if (id >= next_prim_) next_prim_ = id + 1;
    prim_ids_[obj] = id;
    return id;
  }

  int visit(const hittable *h, bool in_boundary, int forced_id = -1) {
    if (!h) throw std::runtime_error("b200rt adaptor: null hittable in scene graph");
    if (auto tag = dynamic_cast<const leaf_tag *>(h)) return visit(tag->inner_.get(), in_boundary, tag->id_);
    if (auto n = dynamic_cast<const bvh_node *>(h)) {
      int me = new_node(B200RT_NODE_BVH);
      int l = visit(n->left_.get(), in_boundary);
      int r = (n->right_.get() == n->left_.get()) ? l : visit(n->right_.get(), in_boundary);
      b200rt_node &d = out.nodes[me];
      d.a = l;
      d.b = r;
      d.f[0] = n->box_.x_.min;
      d.f[1] = n->box_.x_.max;
      d.f[2] = n->box_.y_.min;
      d.f[3] = n->box_.y_.max;
      d.f[4] = n->box_.z_.min;
      d.f[5] = n->box_.z_.max;
      return me;
    }
    if (auto l = dynamic_cast<const hittable_list *>(h)) {
      int me = new_node(B200RT_NODE_LIST);
      std::vector<uint32_t> kids;
      for (const auto &o : l->objects) kids.push_back((uint32_t)visit(o.get(), in_boundary));
      out.nodes[me].a = (int32_t)out.list_children.size();
      out.nodes[me].b = (int32_t)kids.size();
      out.list_children.insert(out.list_children.end(), kids.begin(), kids.end());
      return me;
    }
    if (auto t = dynamic_cast<const translate *>(h)) {
      int me = new_node(B200RT_NODE_TRANSLATE);
      int c = visit(t->object_.get(), in_boundary);
      out.nodes[me].a = c;
      for (int k = 0; k < 3; k++) out.nodes[me].f[k] = t->offset_.e[k];
      return me;
    }
    if (auto r = dynamic_cast<const rotate_x *>(h)) return rotation(B200RT_NODE_ROTATE_X, r->object.get(), r->sin_theta, r->cos_theta, in_boundary);
    if (auto r = dynamic_cast<const rotate_y *>(h)) return rotation(B200RT_NODE_ROTATE_Y, r->object.get(), r->sin_theta, r->cos_theta, in_boundary);
    if (auto r = dynamic_cast<const rotate_z *>(h)) return rotation(B200RT_NODE_ROTATE_Z, r->object.get(), r->sin_theta, r->cos_theta, in_boundary);
    if (auto v = dynamic_cast<const volumne *>(h)) {
      if (in_boundary) throw std::runtime_error("b200rt adaptor: volumne nested inside a volumne boundary is unsupported");
      int me = new_node(B200RT_NODE_VOLUME);
      int id = prim_id(v, forced_id, false);
      b200rt_volume vol;
      vol.boundary_node = visit(v->boundary_.get(), true);
      vol.material = material_row(v->phase_function_.get());
      vol.density = v->density_;
      out.volumes.push_back(vol);
      out.nodes[me].a = (int32_t)out.volumes.size() - 1;
      out.nodes[me].c = id;
      return me;
    }
    if (auto s = dynamic_cast<const sphere *>(h)) {
      int me = new_node(B200RT_NODE_SPHERE);
      out.nodes[me].a = sphere_row(s);
      out.nodes[me].c = prim_id(s, forced_id, in_boundary);
      return me;
    }
    if (auto q = dynamic_cast<const quad *>(h)) {
      int me = new_node(B200RT_NODE_QUAD);
      out.nodes[me].a = quad_row(q);
      out.nodes[me].c = prim_id(q, forced_id, in_boundary);
      return me;
    }
    if (auto t = dynamic_cast<const triangle *>(h)) {
      int me = new_node(B200RT_NODE_TRIANGLE);
      out.nodes[me].a = tri_row(t);
      out.nodes[me].c = prim_id(t, forced_id, in_boundary);
      return me;
    }
    throw std::runtime_error(std::string("b200rt adaptor: unsupported hittable subclass ") + typeid(*h).name());
  }

  int rotation(int kind, const hittable *child, double s, double c, bool in_boundary) {
    int me = new_node(kind);
    int ch = visit(child, in_boundary);
    out.nodes[me].a = ch;
    out.nodes[me].f[0] = s; // values as computed by the reference ctor on the host libm (src/hittable.h:97-98)
    out.nodes[me].f[1] = c;
    return me;
  }

  int sphere_row(const sphere *s) {
    auto it = sphere_rows_.find(s);
    if (it != sphere_rows_.end()) return it->second;
    int row = (int)out.sphere_mat.size();
    put3(out.spheres, s->center_);
    out.spheres.push_back(s->radius_);
    out.spheres.push_back(s->is_moving_ ? 1.0 : 0.0);
    put3(out.spheres, s->begin_center_);
    put3(out.spheres, s->end_center_);
    out.spheres.push_back(0.0);
    out.sphere_mat.push_back(material_row(s->mat_.get()));
    sphere_rows_[s] = row;
    return row;
  }
  int quad_row(const quad *q) {
    auto it = quad_rows_.find(q);
    if (it != quad_rows_.end()) return it->second;
    int row = (int)out.quad_mat.size();
    // the two loop-invariant values quad::hit recomputes per call (src/quad.h:32,39-40), same operations
    double D = dot(q->unorm_, q->corner_);
    vec3 n = cross(q->u_, q->v_);
    vec3 w = n / dot(n, n);
    put3(out.quads, q->corner_);
    put3(out.quads, q->u_);
    put3(out.quads, q->v_);
    put3(out.quads, q->unorm_);
    put3(out.quads, w);
    out.quads.push_back(D);
    out.quads.push_back(q->area_);
    for (int k = 0; k < 3; k++) out.quads.push_back(0.0);
    out.quad_mat.push_back(material_row(q->mat_.get()));
    quad_rows_[q] = row;
    return row;
  }
  int tri_row(const triangle *t) {
    auto it = tri_rows_.find(t);
    if (it != tri_rows_.end()) return it->second;
    int row = (int)out.tri_mat.size();
    put3(out.tris, t->p0_);
    put3(out.tris, t->p1_);
    put3(out.tris, t->p2_);
    put3(out.tris, t->normal_);
    out.tri_mat.push_back(material_row(t->mat_.get()));
    tri_rows_[t] = row;
    return row;
  }

  int material_row(const material *m) {
    if (!m) return -1;
    auto it = mat_rows_.find(m);
    if (it != mat_rows_.end()) return it->second;
    b200rt_material d;
    d.p0 = d.p1 = 0.0f;
    if (auto x = dynamic_cast<const lambertian *>(m)) {
      d.kind = B200RT_MAT_LAMBERTIAN;
      d.tex = texture_row(x->tex_.get());
    } else if (auto x = dynamic_cast<const metal *>(m)) {
      d.kind = B200RT_MAT_METAL;
      d.tex = texture_row(x->tex_.get());
      d.p0 = x->fuzz_;
    } else if (auto x = dynamic_cast<const dielectric *>(m)) {
      d.kind = B200RT_MAT_DIELECTRIC;
      d.tex = texture_row(x->tex_.get());
      d.p0 = x->refract_;
    } else if (auto x = dynamic_cast<const gloss *>(m)) {
      d.kind = B200RT_MAT_GLOSS;
      d.tex = texture_row(x->albedo_.get());
      d.p0 = x->smoothness_;
      d.p1 = x->specular_prob_;
    } else if (auto x = dynamic_cast<const isotropic *>(m)) {
      d.kind = B200RT_MAT_ISOTROPIC;
      d.tex = texture_row(x->tex_.get());
    } else if (auto x = dynamic_cast<const diffuse_light *>(m)) {
      d.kind = B200RT_MAT_DIFFUSE_LIGHT;
      d.tex = texture_row(x->tex_.get());
    } else {
      throw std::runtime_error(std::string("b200rt adaptor: unsupported material subclass ") + typeid(*m).name());
    }
    // glTF scenes give every triangle its own but identical material (src/main.cc:382-383,474-475)
    auto key = std::make_tuple(d.kind, d.tex, d.p0, d.p1);
    auto bv = mat_by_value_.find(key);
    int row;
    if (bv != mat_by_value_.end()) {
      row = bv->second;
    } else {
      row = (int)out.materials.size();
      out.materials.push_back(d);
      mat_by_value_[key] = row;
    }
    mat_rows_[m] = row;
    return row;
  }

  int texture_row(const texture *t) {
    if (!t) throw std::runtime_error("b200rt adaptor: material with a null texture");
    auto it = tex_rows_.find(t);
    if (it != tex_rows_.end()) return it->second;
    b200rt_texture d;
    std::memset(&d, 0, sizeof d);
    d.aux = -1;
    bool by_value = false;
    std::tuple<double, double, double> key;
    if (auto x = dynamic_cast<const solid_color *>(t)) {
      d.kind = B200RT_TEX_SOLID;
      for (int k = 0; k < 3; k++) d.c0[k] = x->color_.e[k];
      key = std::make_tuple(d.c0[0], d.c0[1], d.c0[2]);
      by_value = true;
    } else if (auto x = dynamic_cast<const checker_texture *>(t)) {
      d.kind = B200RT_TEX_CHECKER;
      for (int k = 0; k < 3; k++) d.c0[k] = x->odd_.e[k], d.c1[k] = x->even_.e[k];
      d.scale = x->scale_;
    } else if (auto x = dynamic_cast<const picture_texture *>(t)) {
      d.kind = bilinear_pictures ? B200RT_TEX_PICTURE_LINEAR : B200RT_TEX_PICTURE;
      d.aux = image_row(x->image_.get());
    } else if (auto x = dynamic_cast<const perlin_texture *>(t)) {
      d.kind = B200RT_TEX_PERLIN;
      d.scale = x->scale_;
      d.aux = (int32_t)(out.perlin_perm.size() / 256);
      for (int i = 0; i < 256; i++) {
        put3(out.perlin_vec, x->noise_.rand_offset[i]);
        out.perlin_perm.push_back(x->noise_.perm_x[i]);
      }
    } else if (auto x = dynamic_cast<const value_texture *>(t)) {
      d.kind = B200RT_TEX_VALUE;
      b200rt_value_noise vn;
      vn.resolution = x->noise_.resolution_;
      vn.pad = 0;
      vn.offset = out.value_lattice.size();
      out.value_lattice.insert(out.value_lattice.end(), x->noise_.rand_offset.begin(), x->noise_.rand_offset.end());
      d.aux = (int32_t)out.value_noise.size();
      out.value_noise.push_back(vn);
    } else if (dynamic_cast<const worley_texture *>(t)) {
      d.kind = B200RT_TEX_WORLEY;
    } else if (dynamic_cast<const voronoi_texture *>(t)) {
      d.kind = B200RT_TEX_VORONOI;
    } else {
      throw std::runtime_error(std::string("b200rt adaptor: unsupported texture subclass ") + typeid(*t).name());
    }
    int row;
    if (by_value && solid_by_value_.count(key)) {
      row = solid_by_value_[key];
    } else {
      row = (int)out.textures.size();
      out.textures.push_back(d);
      if (by_value) solid_by_value_[key] = row;
    }
    tex_rows_[t] = row;
    return row;
  }

  int image_row(const image *im) {
    if (!im) throw std::runtime_error("b200rt adaptor: picture_texture with a null image");
    auto it = image_rows_.find(im);
    if (it != image_rows_.end()) return it->second;
    b200rt_image d;
    std::memset(&d, 0, sizeof d);
    d.has_data = (im->bdata != nullptr && im->fdata != nullptr) ? 1 : 0;
    if (d.has_data && float_images) {
      // opt-in: the image as the decoder delivered it, image::fdata (src/image.h:85-86), binary32 per channel,
      // instead of the 8-bit copy the reference samples (src/image.h:99-117): an HDR environment keeps its values above 1
      d.width = im->image_width;
      d.height = im->image_height;
      d.bytes_per_pixel = im->bytes_per_pixel * 4; // 3 channels (stb) -> 12, 4 channels (EXR) -> 16
      d.offset = out.image_bytes.size();
      size_t n = (size_t)im->image_width * im->image_height * im->bytes_per_pixel * sizeof(float);
      const unsigned char *src = reinterpret_cast<const unsigned char *>(im->fdata);
      out.image_bytes.insert(out.image_bytes.end(), src, src + n);
      while (out.image_bytes.size() % 16) out.image_bytes.push_back(0);
    } else if (d.has_data) {
      d.width = im->image_width;
      d.height = im->image_height;
      d.bytes_per_pixel = im->bytes_per_pixel;
      d.offset = out.image_bytes.size();
      size_t n = (size_t)im->image_width * im->image_height * im->bytes_per_pixel;
      out.image_bytes.insert(out.image_bytes.end(), im->bdata, im->bdata + n);
      while (out.image_bytes.size() % 16) out.image_bytes.push_back(0);
    }
    int row = (int)out.images.size();
    out.images.push_back(d);
    image_rows_[im] = row;
    return row;
  }
};

inline b200rt_camera flatten_camera(const camera &cam) {
  b200rt_camera c;
  std::memset(&c, 0, sizeof c);
  c.mode = (int32_t)cam.mode_; // same enumerator order as b200rt_camera_mode (src/camera.h:18)
  c.image_width = cam.image_width_;
  c.image_height = cam.image_height_;
  c.samples_per_pixel = cam.samples_per_pixel_;
  c.max_depth = cam.max_recur_depth_;
  c.viewport_width = cam.viewport_width_;
  c.viewport_height = cam.viewport_height_;
  c.focal_length = cam.mode_ == kLens ? 0.0 : cam.focal_length_; // never read in lens mode (src/camera.h:276-283)
  c.focus_dist = cam.focus_dist_;
  for (int k = 0; k < 3; k++) {
    c.pos[k] = cam.pos_.e[k];
    c.dir[k] = cam.dir_.e[k];
    c.right[k] = cam.right_.e[k];
    c.up[k] = cam.up_.e[k];
    c.defocus_disk_u[k] = cam.defocus_disk_u.e[k];
    c.defocus_disk_v[k] = cam.defocus_disk_v.e[k];
  }
  return c;
}

#ifndef B200RT_ADAPTOR_NO_RENDER
// Drop-in for camera::render (src/camera.h:135-176): same arguments, same P3 output, same cam.image_ contents; the sample
// loop runs on the GPUs listed in `devices` (CUDA ordinals): one device = b200rt_render; several = the frame split by sample
// range across them with one NCCL reduce (b200rt_group_*, include/b200rt.h).  Throws on any backend error.
inline void render(camera &cam, std::ofstream &out, const hittable &world, const std::shared_ptr<const hittable> light,
                   const std::vector<int> &devices, uint64_t seed = 0x5EEDull) {
  if (!out) {
    std::cerr << "Fail to open file.\n"; // src/camera.h:144-147
    return;
  }
  if (devices.empty()) throw std::runtime_error("b200rt: no device given");
  flattener fl;
  fl.run(world, light.get(), cam.background_.get());
  b200rt_scene_desc sd = fl.out.desc();
  b200rt_camera cd = flatten_camera(cam);
  b200rt_group *grp = nullptr;
  auto check = [&](int rc) {
    if (rc != B200RT_OK) {
      std::string msg = b200rt_last_error();
      if (grp) b200rt_group_destroy(grp);
      throw std::runtime_error("b200rt: " + msg);
    }
  };
  check(b200rt_group_create(devices.data(), (int)devices.size(), &grp));
  check(b200rt_group_upload_scene(grp, &sd));
  check(b200rt_group_set_camera(grp, &cd));
  std::vector<double> sum((size_t)cd.image_width * cd.image_height * 3);
  check(b200rt_group_render(grp, seed, 0, cd.samples_per_pixel, sum.data(), nullptr));
  // output stage on the first device: the header and one write_color line per pixel (src/camera.h:149-151, 174; src/color.h:22-36)
  b200rt_ctx *ctx = nullptr;
  check(b200rt_group_context(grp, 0, &ctx));
  std::vector<char> text(b200rt_ppm_bound(cd.image_width, cd.image_height));
  size_t text_len = 0;
  check(b200rt_format_ppm(ctx, sum.data(), cd.samples_per_pixel, text.data(), text.size(), &text_len));
  check(b200rt_finalize(ctx, sum.data(), cd.samples_per_pixel, sum.data(), nullptr));
  b200rt_group_destroy(grp);
  out.write(text.data(), (std::streamsize)text_len);
  cam.image_.resize((size_t)cd.image_width * cd.image_height); // pixel_color / spp, as camera::render leaves it (src/camera.h:169-170)
  for (size_t i = 0; i < cam.image_.size(); i++) cam.image_[i] = color(sum[3 * i], sum[3 * i + 1], sum[3 * i + 2]);
}
inline void render(camera &cam, std::ofstream &out, const hittable &world, const std::shared_ptr<const hittable> light = nullptr,
                   int device = 0, uint64_t seed = 0x5EEDull) {
  render(cam, out, world, light, std::vector<int>{device}, seed);
}
#endif

} // namespace b200
