// scene_file.h — `.akari` scene file -> the flattened akb_scene_desc / akb_render_params of include/akari_b200.h.
//
// Replaces, for the Path hot path, the reference's scene-graph node layer: node registration by name
// (core/nodes/scenegraph.cpp:45-60), SceneNode / PerspectiveCameraNode / AkariMesh / the material nodes /
// PathIntegratorNode / AOIntegratorNode `object_field` + `compile` (core/nodes/{scene,camera,mesh,material,
// integrator}.cpp) and the `.mesh` reader (core/mesh.cpp:48-85). Host-only C++17; owns every array the description
// points into.
#pragma once
#include <memory>
#include <string>
#include <vector>

#include "../../include/akari_b200.h"

namespace akb_host {

struct MeshData {
    std::vector<float> vertices, normals, texcoords;
    std::vector<int32_t> indices, material_indices, material_remap;
};

struct SceneFile {
    std::vector<std::unique_ptr<MeshData>> meshes;
    std::vector<akb_mesh> mesh_descs;
    std::vector<akb_material> materials;
    std::vector<std::unique_ptr<std::vector<float>>> texels; // RGBA float per image texture
    std::vector<akb_texture> textures;
    akb_scene_desc desc;
    akb_render_params params;
    std::string output = "out.png";
    int tile_size = 256; // Path.tile_size / AO.tile_size: the reference's queue/tile extent; informational here
};

// Throws ParseError / std::runtime_error with a file:line:col message.
std::unique_ptr<SceneFile> load_scene_file(const std::string &path);

// core/mesh.cpp:48-85
void read_mesh_file(const std::string &path, MeshData *out);

// 8-bit PNG -> tightly packed RGB bytes, the way `stbi_load(path, &x, &y, &channel, 3)` returns them
// (core/image.cpp:105). png_read.cpp.
void read_png_rgb8(const std::string &path, int *width, int *height, std::vector<uint8_t> *rgb);
// JPEG (baseline, extended sequential and progressive Huffman files) -> the same bytes, bit for bit what stb_image
// v2.23 returns for stbi_load(..., 3). stb picks the decoder by content, so does the front-end. jpeg_read.cpp.
bool looks_like_jpeg(const uint8_t *bytes, size_t n);
void decode_jpeg_rgb8(const std::string &name, const uint8_t *bytes, size_t n, int *width, int *height, std::vector<uint8_t> *rgb);
// Any 8-bit texture file the front-end reads (JPEG or PNG, by content)
void read_image_rgb8(const std::string &path, int *width, int *height, std::vector<uint8_t> *rgb);
// Radiance .hdr -> RGB floats, the way `stbi_loadf(path, &x, &y, &channel, 3)` returns them (core/image.cpp:85-103).
// hdr_read.cpp.
void read_hdr_rgb(const std::string &path, int *width, int *height, std::vector<float> *rgb);

// core/image.cpp:38-61 writes 8-bit PNG through stb; here: a self-contained PNG encoder (stored deflate blocks) and a
// float PFM dump of radiance/weight (the reference has no float output).
void write_png_rgb8(const std::string &path, const uint8_t *rgb, int width, int height);
void write_pfm_rgb(const std::string &path, const float *rgbw, int width, int height);

} // namespace akb_host
