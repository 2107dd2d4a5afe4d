// texpack.cpp -- material / texture table as the path tracer sees it, host side.
//
// Replaces MaterialLoader::load and its helpers (cuda_opengl/src/scene/material_loader.cpp:15-74,
// 164-197, 243-401): from the MTL's per-material constants and the DECODED images (stbi_loadf stays
// the caller's, it is vendored third-party code) it builds
//   * one RGBA float texture per material: rgb = diffuse, a = specular mix (packed from a 3-channel
//     and a 1-channel map, or from one map and the material's constant, or a 1x1 unit texture),
//   * the optional RGB normal map,
//   * Material{diffuse_spec_map, normal_map, ior} with the reference's texture ids.
// The id rules are part of the contract (scene files index into ONE table that grows across scenes):
//   - ids are handed out in material order from a counter that is never reset (the reference's
//     MaterialLoader is a singleton whose _id / caches survive set(), :153-162);
//   - a material with BOTH maps is cached under the concatenated names (:325-327, 393-396): the
//     next material naming the same pair gets the same id;
//   - a material with only ONE map is packed again for every material that names it (:270-306);
//   - a normal map is cached under its own name in the same cache (:399-415); one that failed to
//     load gives -1 (no normal map), a diffuse / specular map that failed gives the unit texture.
// New relative to the reference: the packed texels are written once (the reference packs into a
// new[] buffer, then uploadTextures copies it again), big images are packed by several threads.
#include <cstdlib>
#include <cstring>
#include <memory>
#include <string>
#include <thread>
#include <unordered_map>
#include <vector>

#include "ptcore.h"

struct ptc_texpack
{
  struct Tex
  {
    int32_t w = 0, h = 0, nb_chan = 0;
    std::shared_ptr<std::vector<float>> own; // packed / unit / copied texels
    const float* data = nullptr;
  };
  std::vector<Tex> textures;                         // MaterialLoader::_textures
  std::unordered_map<std::string, int> packed;       // _packed_tex: token -> id
  std::unordered_map<std::string, Tex> loaded;       // _loaded_tex: name -> decoded image (copied)
  unsigned id = 0;                                   // _id
  std::string err;
};

namespace {

using Tex = ptc_texpack::Tex;

template <typename F>
void
parallel_pixels(size_t n, F f)
{
  const unsigned hw = std::thread::hardware_concurrency();
  const unsigned t = n >= (1u << 18) ? (hw ? (hw < 16u ? hw : 16u) : 1u) : 1u;
  if (t <= 1) {
    f((size_t)0, n);
    return;
  }
  std::vector<std::thread> th;
  for (unsigned k = 1; k < t; ++k)
    th.emplace_back([=, &f] { f(n * k / t, n * (k + 1) / t); });
  f((size_t)0, n / t);
  for (std::thread& x : th)
    x.join();
}

Tex
make_tex(int32_t w, int32_t h, int32_t nb_chan)
{
  Tex t;
  t.w = w, t.h = h, t.nb_chan = nb_chan;
  t.own = std::make_shared<std::vector<float>>((size_t)w * h * nb_chan);
  t.data = t.own->data();
  return t;
}

// createUnitTex(float4), material_loader.cpp:139-149
Tex
unit_tex(const float rgb[3], float a)
{
  Tex t = make_tex(1, 1, 4);
  float* d = t.own->data();
  d[0] = rgb[0], d[1] = rgb[1], d[2] = rgb[2], d[3] = a;
  return t;
}

// pack(rgb, a), :15-36 -- sizes already equal
Tex
pack_rgb_a(const Tex& rgb, const Tex& a)
{
  Tex r = make_tex(rgb.w, rgb.h, 4);
  float* d = r.own->data();
  parallel_pixels((size_t)rgb.w * rgb.h, [&](size_t lo, size_t hi) {
    for (size_t p = lo; p < hi; ++p) {
      d[4 * p] = rgb.data[3 * p], d[4 * p + 1] = rgb.data[3 * p + 1], d[4 * p + 2] = rgb.data[3 * p + 2];
      d[4 * p + 3] = a.data[p];
    }
  });
  return r;
}

// pack(tex, float3 default), :38-52 -- a 1-channel map under a constant colour
Tex
pack_const_a(const Tex& a, const float rgb[3])
{
  Tex r = make_tex(a.w, a.h, 4);
  float* d = r.own->data();
  parallel_pixels((size_t)a.w * a.h, [&](size_t lo, size_t hi) {
    for (size_t p = lo; p < hi; ++p)
      d[4 * p] = rgb[0], d[4 * p + 1] = rgb[1], d[4 * p + 2] = rgb[2], d[4 * p + 3] = a.data[p];
  });
  return r;
}

// pack(tex, float default), :60-74 -- a 3-channel map with a constant specular mix
Tex
pack_rgb_const(const Tex& rgb, float a)
{
  Tex r = make_tex(rgb.w, rgb.h, 4);
  float* d = r.own->data();
  parallel_pixels((size_t)rgb.w * rgb.h, [&](size_t lo, size_t hi) {
    for (size_t p = lo; p < hi; ++p) {
      d[4 * p] = rgb.data[3 * p], d[4 * p + 1] = rgb.data[3 * p + 1], d[4 * p + 2] = rgb.data[3 * p + 2];
      d[4 * p + 3] = a;
    }
  });
  return r;
}

// checkAndupload, :84-114: the first image seen under a name is THE image of that name
const Tex&
load_once(ptc_texpack* p, const ptc_image* img)
{
  const std::string name = img->name ? img->name : "";
  auto it = p->loaded.find(name);
  if (it != p->loaded.end())
    return it->second;
  Tex t;
  if (img->data && img->w > 0 && img->h > 0 && img->nb_chan > 0) {
    t = make_tex(img->w, img->h, img->nb_chan);
    memcpy(t.own->data(), img->data, t.own->size() * sizeof(float));
  } // else: failed to load -> {0, 0, 0, null}
  return p->loaded.emplace(name, std::move(t)).first->second;
}

inline bool
present(const ptc_image* img)
{
  return img && img->name && img->name[0];
}

} // namespace

namespace {

// getTextureId(tex_rgb, tex_a, default_rgb, default_a), material_loader.cpp:243-397
int
diffuse_spec_id(ptc_texpack* p, const ptc_image* rgb_img, const ptc_image* a_img, const float rgb[3], float a,
                int* out)
{
  const bool has_rgb = present(rgb_img), has_a = present(a_img);
  if (!has_rgb && !has_a) { // CASE 1: no map at all -> 1x1 texture of the material's constants
    p->textures.push_back(unit_tex(rgb, a));
    *out = (int)p->id++;
    return PTC_OK;
  }
  const Tex* t_rgb = has_rgb ? &load_once(p, rgb_img) : nullptr;
  const Tex* t_a = has_a ? &load_once(p, a_img) : nullptr;
  if (!has_rgb) { // CASE 2a: specular map only -- packed again for every material that names it
    *out = (int)p->id++;
    p->textures.push_back(t_a->nb_chan == 1 ? pack_const_a(*t_a, rgb) : unit_tex(rgb, a));
    return PTC_OK;
  }
  if (!has_a) { // CASE 2b: diffuse map only
    *out = (int)p->id++;
    p->textures.push_back(t_rgb->nb_chan == 3 ? pack_rgb_const(*t_rgb, a) : unit_tex(rgb, a));
    return PTC_OK;
  }
  // CASE 3: both maps -> packed once, cached under the concatenated names
  const std::string token = std::string(rgb_img->name) + a_img->name;
  auto hit = p->packed.find(token);
  if (hit != p->packed.end()) {
    *out = hit->second;
    return PTC_OK;
  }
  if (t_rgb->nb_chan != 3 || t_a->nb_chan != 1) {
    const int curr = (int)p->id++;
    *out = curr;
    if (t_rgb->nb_chan != 3 && t_a->nb_chan != 1) { // both unusable: unit texture, NOT cached (:335-343)
      p->textures.push_back(unit_tex(rgb, a));
      return PTC_OK;
    }
    p->textures.push_back(t_rgb->nb_chan != 3 ? pack_const_a(*t_a, rgb) : pack_rgb_const(*t_rgb, a));
    p->packed[token] = curr;
    return PTC_OK;
  }
  if (t_rgb->w != t_a->w || t_rgb->h != t_a->h) {
    // the reference rescales the smaller map with stb_image_resize (:367-389); that is third-party
    // code this library does not carry -- the caller resizes (stbir_resize_float) and calls again
    p->err = std::string("diffuse map '") + rgb_img->name + "' and specular map '" + a_img->name +
             "' differ in size: resize the smaller one to the larger before packing";
    return PTC_ERR_INVALID;
  }
  p->textures.push_back(pack_rgb_a(*t_rgb, *t_a));
  *out = p->packed[token] = (int)p->id++;
  return PTC_OK;
}

// getTextureId(tex_rgb) -> registerOrGet, :227-234, 399-415
int
normal_id(ptc_texpack* p, const ptc_image* img)
{
  if (!present(img))
    return -1;
  auto hit = p->packed.find(img->name);
  if (hit != p->packed.end())
    return hit->second;
  const Tex& t = load_once(p, img);
  if (t.nb_chan == 0)
    return -1; // failed to load: the material simply has no normal map
  p->packed[img->name] = (int)p->id;
  p->textures.push_back(t); // the decoded image itself, whatever its channel count
  return (int)p->id++;
}

} // namespace

extern "C" {

int
ptc_texpack_create(ptc_texpack** out)
{
  if (!out)
    return PTC_ERR_INVALID;
  *out = new ptc_texpack();
  return PTC_OK;
}

void
ptc_texpack_destroy(ptc_texpack* p)
{
  delete p;
}

const char*
ptc_texpack_last_error(const ptc_texpack* p)
{
  return p ? p->err.c_str() : "null packer";
}

// MaterialLoader::set + load for the materials of ONE scene (material_loader.cpp:153-197)
int
ptc_texpack_add_materials(ptc_texpack* p, const ptc_mtl_entry* mats, uint32_t n, ptc_material* out)
{
  if (!p || (n && (!mats || !out)))
    return PTC_ERR_INVALID;
  for (uint32_t i = 0; i < n; ++i) {
    const ptc_mtl_entry& m = mats[i];
    // default specular mix = mean of Ks, divided in double like the reference's `/ 3.0` (:174-176)
    const float spec = (float)((double)(m.specular[0] + m.specular[1] + m.specular[2]) / 3.0);
    int id = -1;
    const int rc = diffuse_spec_id(p, m.diffuse_map, m.specular_map, m.diffuse, spec, &id);
    if (rc)
      return rc;
    out[i].diffuse_spec_map = id;
    out[i].normal_map = normal_id(p, m.normal_map);
    out[i].ior = m.ior;
    out[i]._pad = 0;
  }
  return PTC_OK;
}

uint32_t
ptc_texpack_texture_count(const ptc_texpack* p)
{
  return p ? (uint32_t)p->textures.size() : 0u;
}

// The global texture table (Scenes.textures) so far; `data` pointers stay valid until the packer
// is destroyed.  Pass the array as ptc_scene_desc.textures.
int
ptc_texpack_textures(const ptc_texpack* p, ptc_texture* out, uint32_t capacity)
{
  if (!p || (!out && capacity))
    return PTC_ERR_INVALID;
  if (capacity < p->textures.size())
    return PTC_ERR_INVALID;
  for (size_t i = 0; i < p->textures.size(); ++i) {
    memset(&out[i], 0, sizeof(ptc_texture));
    out[i].w = p->textures[i].w, out[i].h = p->textures[i].h, out[i].nb_chan = p->textures[i].nb_chan;
    out[i].data = p->textures[i].data;
  }
  return PTC_OK;
}

} // extern "C"
