// TEST INFRASTRUCTURE (oracle/_ref recipe) -- not product code.
//
// Runs the reference's OWN host loader (scene::Scene::upload,
// cuda_opengl/src/scene/scene.cpp:304-358 and MaterialLoader,
// cuda_opengl/src/scene/material_loader.cpp:153-401), compiled from the
// sources where they lie under /root/reference and linked against
// ref_shim/cuda_host_shim.cpp, then dumps what the path tracer kernel would
// have received (Face[], Material[], LightProp[], Texture[], Camera, cubemap)
// into a ".ptscene" container.  The container is the golden fixture the oracle,
// the tests and the bench consume on machines where /root/reference is absent.
//
// usage: ref_loader ASSET_FOLDER SCENE_FILE OUT.ptscene
#define STB_IMAGE_IMPLEMENTATION
#include <stb/stb_image.h>

#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <iostream>
#include <string>
#include <vector>

#include <scene/material_loader.h>
#include <scene/scene.h>
#include <utils/texture_utils.h>
#include <utils/utils.h>

namespace {

struct Writer
{
  FILE* f;
  void raw(const void* p, size_t n) { fwrite(p, 1, n, f); }
  void u32(uint32_t v) { raw(&v, 4); }
  void i32(int32_t v) { raw(&v, 4); }
  void f32(float v) { raw(&v, 4); }
};

// Same constant-colour cubemap as cuda_opengl/src/gpu_processor.cpp:38-58.
std::vector<float>
unit_cubemap(unsigned color)
{
  float r = ((color >> 16) & 0xFF) / 255.0f;
  float g = ((color >> 8) & 0xFF) / 255.0f;
  float b = (color & 0xFF) / 255.0f;
  std::vector<float> img(4 * 6);
  for (int n = 0; n < 6; ++n) {
    img[n * 4 + 0] = r;
    img[n * 4 + 1] = g;
    img[n * 4 + 2] = b;
    img[n * 4 + 3] = 0.0f;
  }
  return img;
}

// Follows uploadCubemap(), cuda_opengl/src/gpu_processor.cpp:68-161, up to the
// point where the faces are handed to cudaMemcpy3D.
std::vector<float>
build_cubemap(const std::string& path_in_scene, const std::string& folder,
              unsigned& size)
{
  const unsigned DEFAULT_COLOR = 0x131b23;
  // uploadCubemaps() (gpu_processor.cpp:163-182) turns an empty path into the
  // literal "empty" and always prefixes the asset folder.
  std::string path = folder + "/" + (path_in_scene.empty() ? "empty" : path_in_scene);
  size = 1;
  if (path.empty() || utils::isHexa(path))
    return unit_cubemap(path.empty() ? DEFAULT_COLOR
                                     : strtol(path.c_str(), NULL, 16));
  int w = 0, h = 0, nb_chan = 0;
  float* loaded = stbi_loadf(path.c_str(), &w, &h, &nb_chan, STBI_default);
  std::string error;
  unsigned s = loaded ? w / 4 : 0;
  if (!loaded)
    error = "unknown error " + path;
  else if (s != (unsigned)(h) / 3)
    error = "width and height are not the same";
  else if (s & (s - 1))
    error = "size should be a power of 2";
  if (!error.empty()) {
    std::cerr << "ref_loader: cubemap fallback: " << error << std::endl;
    if (loaded)
      stbi_image_free(loaded);
    return unit_cubemap(DEFAULT_COLOR);
  }
  size = s;
  std::vector<float> img((size_t)size * size * 4 * 6);
  float* tmp = texture::append_cube_faces(img.data(), loaded, w, 0, nb_chan, 4,
                                          true, true);
  tmp = texture::append_cube_faces(tmp, loaded, w, 1, nb_chan, 4, false, false);
  texture::append_cube_faces(tmp, loaded, w, 1, nb_chan, 4, true, false);
  stbi_image_free(loaded);
  return img;
}

// stb_image.h:1565-1582 (stbi__ldr_to_hdr): float = (float)pow(u8/255.0f, 2.2f).
inline float
stb_gamma(unsigned char v)
{
  return (float)(pow(v / 255.0f, 2.2f) * 1.0f);
}

} // namespace

int
main(int argc, char** argv)
{
  // ref_loader --cross IN.f32 W H NB_CHAN OUT.f32 : the reference's own face extraction
  // (texture::append_cube_faces in uploadCubemap's order, gpu_processor.cpp:116-127) on a raw
  // float cube cross; writes u32 size + 6*size*size*4 floats.  Used to pin ptc_cubemap_from_cross.
  if (argc == 7 && std::string(argv[1]) == "--cross") {
    const int w = atoi(argv[3]), h = atoi(argv[4]), nb_chan = atoi(argv[5]);
    std::vector<float> in((size_t)w * h * nb_chan);
    FILE* fi = fopen(argv[2], "rb");
    if (!fi || fread(in.data(), 4, in.size(), fi) != in.size()) {
      perror("ref_loader: --cross input");
      return 1;
    }
    fclose(fi);
    const unsigned size = w / 4; // gpu_processor.cpp:101
    std::vector<float> img((size_t)size * size * 4 * 6);
    float* tmp = texture::append_cube_faces(img.data(), in.data(), w, 0, nb_chan, 4, true, true);
    tmp = texture::append_cube_faces(tmp, in.data(), w, 1, nb_chan, 4, false, false);
    texture::append_cube_faces(tmp, in.data(), w, 1, nb_chan, 4, true, false);
    FILE* fo = fopen(argv[6], "wb");
    if (!fo) {
      perror("ref_loader: --cross output");
      return 1;
    }
    fwrite(&size, 4, 1, fo);
    fwrite(img.data(), 4, img.size(), fo);
    fclose(fo);
    return 0;
  }
  // ref_loader --cross-u8 IN.jpg OUT.bin : a shipped cube cross (assets/cubemap/*.jpg) as the 8-bit
  // texels stb_image decodes + the 256 floats stbi_loadf turns them into (uploadCubemap loads with
  // stbi_loadf, gpu_processor.cpp:96; stb_image.h:1565-1582 maps every 8-bit value through one
  // pow()).  Checks here that table[u8] reproduces stbi_loadf's output exactly, so the GPU box can
  // rebuild the float cross from 36 MB instead of 151 MB.  Layout: u32 w, h, nb_chan; 256 floats;
  // w*h*nb_chan bytes.
  if (argc == 4 && std::string(argv[1]) == "--cross-u8") {
    int w = 0, h = 0, nc = 0, w2 = 0, h2 = 0, nc2 = 0;
    unsigned char* u8 = stbi_load(argv[2], &w, &h, &nc, STBI_default);
    float* f32 = stbi_loadf(argv[2], &w2, &h2, &nc2, STBI_default);
    if (!u8 || !f32 || w != w2 || h != h2 || nc != nc2) {
      std::cerr << "ref_loader: --cross-u8: cannot decode " << argv[2] << std::endl;
      return 1;
    }
    float table[256];
    for (int v = 0; v < 256; ++v)
      table[v] = stb_gamma((unsigned char)v);
    const size_t n = (size_t)w * h * nc;
    for (size_t i = 0; i < n; ++i)
      if (memcmp(&table[u8[i]], &f32[i], 4) != 0) {
        std::cerr << "ref_loader: --cross-u8: stbi_loadf differs from the table at texel " << i << std::endl;
        return 1;
      }
    FILE* fo = fopen(argv[3], "wb");
    if (!fo) {
      perror("ref_loader: --cross-u8 output");
      return 1;
    }
    const uint32_t hdr[3] = { (uint32_t)w, (uint32_t)h, (uint32_t)nc };
    fwrite(hdr, 4, 3, fo);
    fwrite(table, 4, 256, fo);
    fwrite(u8, 1, n, fo);
    fclose(fo);
    stbi_image_free(u8);
    stbi_image_free(f32);
    return 0;
  }
  if (argc < 4) {
    std::cerr << "usage: ref_loader ASSET_FOLDER SCENE_FILE OUT.ptscene\n";
    return 2;
  }
  // ref_loader --indexed OBJ_PATH MTL_DIR OUT.ptix : what the reference's vendored tinyobjloader
  // hands to upload_meshes (scene.cpp:202-283) for this OBJ -- the attribute arrays, and per shape
  // the index triples and per-face material ids -- BEFORE the reference flattens it into Face[].
  // Pins ptc_flatten_faces: its output for this file must equal the Face[] of the .ptscene dump.
  if (argc == 5 && std::string(argv[1]) == "--indexed") {
    tinyobj::attrib_t attrib;
    std::vector<tinyobj::shape_t> shapes;
    std::vector<tinyobj::material_t> materials;
    std::string err;
    if (!tinyobj::LoadObj(&attrib, &shapes, &materials, &err, argv[2], argv[3])) { // scene.cpp:341
      std::cerr << "ref_loader: --indexed: " << err << "\n";
      return 1;
    }
    FILE* fo = fopen(argv[4], "wb");
    if (!fo) {
      perror("ref_loader: --indexed output");
      return 1;
    }
    Writer w{ fo };
    w.raw("PTIX", 4);
    w.u32(1);
    auto farr = [&](const std::vector<float>& v) {
      uint64_t n = v.size();
      w.raw(&n, 8);
      w.raw(v.data(), n * 4);
    };
    farr(attrib.vertices);
    farr(attrib.normals);
    farr(attrib.texcoords);
    w.u32((uint32_t)shapes.size());
    static_assert(sizeof(tinyobj::index_t) % 4 == 0, "index_t is a record of ints");
    w.u32((uint32_t)(sizeof(tinyobj::index_t) / 4)); // ints per index record (vertex, normal, texcoord first)
    for (const tinyobj::shape_t& sh : shapes) {
      uint64_t ni = sh.mesh.indices.size(), nm = sh.mesh.material_ids.size();
      w.raw(&ni, 8);
      w.raw(sh.mesh.indices.data(), ni * sizeof(tinyobj::index_t));
      w.raw(&nm, 8);
      w.raw(sh.mesh.material_ids.data(), nm * 4);
    }
    fclose(fo);
    return 0;
  }
  // ref_loader --materials OBJ_PATH MTL_DIR OUT.ptmt : the INPUT of MaterialLoader::load
  // (material_loader.cpp:164-197) for this OBJ: per tinyobj material Kd, Ks, ior and the three
  // texture names it looks at (diffuse, specular, bump-or-normal), then every distinct image
  // decoded exactly as checkAndupload does (stbi_loadf(base + "/" + name, STBI_default), :84-114).
  // Pins ptc_texpack_*: its materials / texture table for this input must equal the .ptscene dump.
  if (argc == 5 && std::string(argv[1]) == "--materials") {
    tinyobj::attrib_t attrib;
    std::vector<tinyobj::shape_t> shapes;
    std::vector<tinyobj::material_t> materials;
    std::string err;
    if (!tinyobj::LoadObj(&attrib, &shapes, &materials, &err, argv[2], argv[3])) {
      std::cerr << "ref_loader: --materials: " << err << "\n";
      return 1;
    }
    FILE* fo = fopen(argv[4], "wb");
    if (!fo) {
      perror("ref_loader: --materials output");
      return 1;
    }
    Writer w{ fo };
    w.raw("PTMT", 4);
    w.u32(1);
    auto str = [&](const std::string& s) {
      w.u32((uint32_t)s.size());
      w.raw(s.data(), s.size());
    };
    std::vector<std::string> names;
    auto want = [&](const std::string& n) {
      if (!n.empty() && std::find(names.begin(), names.end(), n) == names.end())
        names.push_back(n);
    };
    w.u32((uint32_t)materials.size());
    for (const tinyobj::material_t& m : materials) {
      for (int k = 0; k < 3; ++k)
        w.f32(m.diffuse[k]);
      for (int k = 0; k < 3; ++k)
        w.f32(m.specular[k]);
      w.f32(m.ior);
      const std::string nm = !m.bump_texname.empty() ? m.bump_texname : m.normal_texname; // :184-187
      str(m.diffuse_texname), str(m.specular_texname), str(nm);
      want(m.diffuse_texname), want(m.specular_texname), want(nm);
    }
    w.u32((uint32_t)names.size());
    for (const std::string& n : names) {
      int iw = 0, ih = 0, nc = 0;
      float* px = stbi_loadf((std::string(argv[3]) + "/" + n).c_str(), &iw, &ih, &nc, STBI_default);
      str(n);
      if (!px)
        iw = ih = nc = 0;
      w.i32(iw), w.i32(ih), w.i32(nc);
      if (px) {
        w.raw(px, (size_t)iw * ih * nc * 4);
        stbi_image_free(px);
      }
    }
    fclose(fo);
    return 0;
  }
  std::string folder = argv[1], scene_file = argv[2], out = argv[3];

  scene::Scene sc(folder + "/" + scene_file); // gpu_processor.cpp:262-263
  sc.upload(nullptr);                          // gpu_processor.cpp:289
  if (!sc.ready()) {
    std::cerr << "ref_loader: scene failed to load: " << sc.error() << "\n";
    return 1;
  }
  const scene::SceneData* sd = sc.getUploadedScenePointer(); // host memory (shim)
  const scene::Camera cam = sc.getInitCamera();
  const auto& textures = scene::MaterialLoader::instance()->getTextures();

  FILE* f = fopen(out.c_str(), "wb");
  if (!f) {
    perror("ref_loader: fopen");
    return 1;
  }
  Writer w{ f };
  w.raw("PTSC", 4);
  w.u32(1);
  static_assert(sizeof(scene::Camera) == 64, "Camera layout");
  static_assert(sizeof(scene::Face) == 112, "Face layout");
  static_assert(sizeof(scene::LightProp) == 32, "LightProp layout");
  w.raw(&cam, sizeof(cam));

  w.u32(sd->lights.size);
  for (unsigned i = 0; i < sd->lights.size; ++i)
    w.raw(&sd->lights.data[i], sizeof(scene::LightProp));

  w.u32(sd->materials.size);
  for (unsigned i = 0; i < sd->materials.size; ++i) {
    const scene::Material& m = sd->materials.data[i];
    w.i32(m.diffuse_spec_map);
    w.i32(m.normal_map);
    w.f32(m.ior);
  }

  w.u32(sd->meshes.size);
  size_t total_faces = 0;
  for (unsigned i = 0; i < sd->meshes.size; ++i) {
    w.u32(sd->meshes.data[i].faces.size);
    total_faces += sd->meshes.data[i].faces.size;
  }
  for (unsigned i = 0; i < sd->meshes.size; ++i) {
    const auto& faces = sd->meshes.data[i].faces;
    if (faces.size)
      w.raw(faces.data, (size_t)faces.size * sizeof(scene::Face));
  }

  w.u32((uint32_t)textures.size());
  size_t tex_float_bytes = 0;
  for (const scene::Texture& t : textures) {
    w.i32(t.w);
    w.i32(t.h);
    w.i32(t.nb_chan);
    size_t n = (size_t)t.w * t.h * t.nb_chan;
    tex_float_bytes += n * 4;
    // Try the compact form: every channel is an stb gamma-decoded byte, the
    // alpha channel possibly a constant (pack(tex, default_a),
    // material_loader.cpp:59-74).
    bool compact = n > 64;
    bool alpha_const = false;
    float alpha_v = 0.f;
    std::vector<unsigned char> bytes;
    if (compact) {
      int nc = t.nb_chan;
      if (nc == 4) {
        alpha_const = true;
        alpha_v = t.data[3];
        for (size_t i = 3; i < n; i += 4)
          if (t.data[i] != alpha_v) {
            alpha_const = false;
            break;
          }
      }
      int stored = (nc == 4 && alpha_const) ? 3 : nc;
      bytes.resize((size_t)t.w * t.h * stored);
      size_t k = 0;
      for (size_t px = 0; px < (size_t)t.w * t.h && compact; ++px)
        for (int c = 0; c < stored; ++c) {
          float v = t.data[px * nc + c];
          long q = lround(255.0 * pow((double)v, 1.0 / 2.2));
          if (q < 0 || q > 255 || stb_gamma((unsigned char)q) != v) {
            compact = false;
            break;
          }
          bytes[k++] = (unsigned char)q;
        }
    }
    if (compact) {
      w.u32(1);
      w.u32(alpha_const ? 1 : 0);
      w.f32(alpha_v);
      w.raw(bytes.data(), bytes.size());
    } else {
      w.u32(0);
      w.raw(t.data, n * 4);
    }
  }

  unsigned cube_size = 1;
  std::vector<float> cube = build_cubemap(sc.getCubemapPath(), folder, cube_size);
  const std::string& cp = sc.getCubemapPath();
  w.u32((uint32_t)cp.size());
  w.raw(cp.data(), cp.size());
  w.u32(cube_size);
  if (cube_size <= 64) {
    w.u32(1);
    w.raw(cube.data(), cube.size() * 4);
  } else {
    w.u32(0); // too large for a fixture; faces omitted
  }
  fclose(f);

  std::cout << "ref_loader: " << scene_file << ": meshes=" << sd->meshes.size
            << " faces=" << total_faces << " materials=" << sd->materials.size
            << " lights=" << sd->lights.size << " textures=" << textures.size()
            << " (" << tex_float_bytes << " B as float) cubemap=" << cube_size
            << "^2x6 -> " << out << std::endl;
  return 0;
}
