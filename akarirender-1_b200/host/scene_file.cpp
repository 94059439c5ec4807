// scene_file.cpp — see scene_file.h. Host arithmetic follows the reference's node code operation for operation
// (compiled with -ffp-contract=off): a scene file must flatten to the very floats the reference's own nodes produce.
#include "scene_file.h"

#include <math.h>
#include <stdio.h>
#include <string.h>

#include <algorithm>
#include <map>

#include "sdl.h"

namespace akb_host {
namespace {

const char *const kNodeTypes[] = {"Scene",          "PerspectiveCamera", "AkariMesh",   "DiffuseMaterial",
                                  "GlossyMaterial", "EmissiveMaterial",  "MixMaterial", "Path",
                                  "AO",
                                  // extension, not a reference node:
                                  "DielectricMaterial"};

bool known_type(const std::string &t) {
    for (const char *k : kNodeTypes)
        if (t == k) return true;
    return false;
}

[[noreturn]] void bad(const Object &o, const std::string &msg) { throw std::runtime_error(o.file + ": " + o.type + ": " + msg); }

double num(const Object &o, const char *field, const Value &v) {
    if (!v.is_number()) bad(o, std::string("field ") + field + " must be a number");
    return v.num;
}

// load_array<floatN> (core/nodes/scenegraph.h:75-84): each element is read as double and narrowed
void float_array(const Object &o, const char *field, const Value &v, float *out, int n) {
    if (!v.is_array() || (int)v.arr.size() != n) bad(o, std::string("field ") + field + " must be an array of " + std::to_string(n) + " numbers");
    for (int i = 0; i < n; i++) out[i] = (float)num(o, field, v.arr[i]);
}

// ---- 4x4 helpers in the reference's evaluation order (common/math.h:97-105, :266-293)
struct M4 {
    float m[4][4];
};
M4 identity() {
    M4 r;
    memset(&r, 0, sizeof r);
    for (int i = 0; i < 4; i++) r.m[i][i] = 1.0f;
    return r;
}
M4 mul(const M4 &a, const M4 &b) {
    M4 r;
    for (int i = 0; i < 4; i++)
        for (int j = 0; j < 4; j++) {
            float s = a.m[i][0] * b.m[0][j];
            for (int k = 1; k < 4; k++) s += a.m[i][k] * b.m[k][j];
            r.m[i][j] = s;
        }
    return r;
}
M4 rotate(int axis, float theta) {
    const float s = sinf(theta), c = cosf(theta);
    M4 r = identity();
    if (axis == 0)
        r.m[1][1] = c, r.m[1][2] = -s, r.m[2][1] = s, r.m[2][2] = c;
    else if (axis == 1)
        r.m[0][0] = c, r.m[0][2] = s, r.m[2][0] = -s, r.m[2][2] = c;
    else
        r.m[0][0] = c, r.m[0][1] = -s, r.m[1][0] = s, r.m[1][1] = c;
    return r;
}

const float kPiF = 3.1415926535897932384f; // Constants<Float>::Pi(), common/math.h:37 (a float literal)

// PerspectiveCameraNode::object_field + compile, core/nodes/camera.cpp:26-51
akb_camera camera_from(const Object &o) {
    float position[3] = {0, 0, 0}, rotation[3] = {0, 0, 0};
    int res[2] = {512, 512};
    double fov = (double)(80.0f * kPiF / 180.0f); // `double fov = radians(80.0f)`: evaluated in float
    for (auto &f : o.fields) {
        if (f.first == "fov") {
            fov = num(o, "fov", f.second) * (double)kPiF / 180.0f; // radians<double>: Pi() is still the float literal
        } else if (f.first == "rotation") {
            float deg[3];
            float_array(o, "rotation", f.second, deg, 3);
            for (int i = 0; i < 3; i++) rotation[i] = deg[i] * kPiF / 180.0f;
        } else if (f.first == "position") {
            float_array(o, "position", f.second, position, 3);
        } else if (f.first == "resolution") {
            if (!f.second.is_array() || f.second.arr.size() != 2) bad(o, "field resolution must be [width, height]");
            for (int i = 0; i < 2; i++) res[i] = (int)num(o, "resolution", f.second.arr[i]);
        }
    }
    M4 c2w = rotate(2, rotation[2]);
    c2w = mul(rotate(0, rotation[1]), c2w);
    c2w = mul(rotate(1, rotation[0]), c2w);
    M4 t = identity();
    t.m[0][3] = position[0], t.m[1][3] = position[1], t.m[2][3] = position[2];
    c2w = mul(t, c2w);
    akb_camera cam;
    memset(&cam, 0, sizeof cam);
    memcpy(cam.c2w, c2w.m, sizeof cam.c2w);
    cam.fov = (float)fov; // PerspectiveCamera(resolution, c2w, Float fov)
    cam.width = res[0], cam.height = res[1];
    return cam;
}

struct Flattener {
    SceneFile *sf;
    std::map<const Object *, int> index; // scene-wide material table, one entry per distinct node
    std::map<std::string, int> texture_of; // image textures by resolved path

    // ImageTextureNode + DefaultImageReader (core/nodes/material.cpp:46-71, core/image.cpp:78-125): `.hdr` files are
    // float RGB as stored (stbi_loadf), everything else bytes / 255.0f (stbi_load), alpha 1. The reference opens the
    // path relative to the directory of the file being parsed.
    int image_texture(const Object &o, const std::string &rel) {
        const std::string path = join_path(dir_of(o.file), rel);
        auto it = texture_of.find(path);
        if (it != texture_of.end()) return it->second;
        int w = 0, h = 0;
        std::unique_ptr<std::vector<float>> tex;
        const size_t dot = path.rfind('.');
        if (dot != std::string::npos && path.substr(dot) == ".hdr") { // the reference decides by extension, exactly ".hdr"
            std::vector<float> rgb;
            read_hdr_rgb(path, &w, &h, &rgb);
            tex = std::make_unique<std::vector<float>>((size_t)w * h * 4);
            for (size_t i = 0; i < (size_t)w * h; i++) {
                for (int k = 0; k < 3; k++) (*tex)[4 * i + k] = rgb[3 * i + k];
                (*tex)[4 * i + 3] = 1.0f;
            }
        } else {
            std::vector<uint8_t> rgb;
            read_image_rgb8(path, &w, &h, &rgb);
            tex = std::make_unique<std::vector<float>>((size_t)w * h * 4);
            for (size_t i = 0; i < (size_t)w * h; i++) {
                for (int k = 0; k < 3; k++) (*tex)[4 * i + k] = (float)rgb[3 * i + k] / 255.0f;
                (*tex)[4 * i + 3] = 1.0f;
            }
        }
        akb_texture d;
        d.width = w, d.height = h, d.rgba = tex->data();
        const int id = (int)sf->textures.size();
        sf->textures.push_back(d);
        sf->texels.push_back(std::move(tex));
        texture_of[path] = id;
        return id;
    }

    // resolve_texture (core/nodes/material.cpp:73-88): [r,g,b] | number | "image path". Returns the texture index
    // (-1 for a constant, whose value goes to rgb).
    int texture_rgb(const Object &o, const char *field, const Value &v, float rgb[3]) {
        rgb[0] = rgb[1] = rgb[2] = 0.0f;
        if (v.is_array()) {
            float_array(o, field, v, rgb, 3);
        } else if (v.is_number()) {
            rgb[0] = rgb[1] = rgb[2] = (float)v.num;
        } else if (v.is_string()) {
            return image_texture(o, v.str);
        } else {
            bad(o, std::string("field ") + field + " must be [r,g,b], a number or an image path");
        }
        return -1;
    }

    const Object &material_node(const Object &owner, const char *field, const Value &v) {
        if (!v.is_object()) bad(owner, std::string("field ") + field + " must be a material node");
        const std::string &t = v.obj->type;
        if (t.size() < 8 || t.compare(t.size() - 8, 8, "Material") != 0) bad(owner, std::string("field ") + field + ": " + t + " is not a material");
        return *v.obj;
    }

    int material_index(const Object &m) {
        auto it = index.find(&m);
        if (it != index.end()) return it->second;
        const int id = (int)sf->materials.size();
        index[&m] = id;
        akb_material e;
        memset(&e, 0, sizeof e);
        e.child_a = e.child_b = -1;
        e.color_tex = e.roughness_tex = e.fraction_tex = -1;
        sf->materials.push_back(e); // children are appended after their parent
        if (m.type == "DiffuseMaterial" || m.type == "EmissiveMaterial") {
            e.type = m.type == "DiffuseMaterial" ? AKB_MAT_DIFFUSE : AKB_MAT_EMISSIVE;
            const Value *c = m.find("color");
            if (!c) bad(m, "field color missing");
            e.color_tex = texture_rgb(m, "color", *c, e.color);
        } else if (m.type == "GlossyMaterial") {
            e.type = AKB_MAT_GLOSSY;
            const Value *c = m.find("color"), *r = m.find("roughness");
            if (!c || !r) bad(m, "fields color and roughness are required");
            e.color_tex = texture_rgb(m, "color", *c, e.color);
            float rr[3];
            e.roughness_tex = texture_rgb(m, "roughness", *r, rr);
            e.roughness = rr[0]; // kernel/material.h:225-226 reads channel 0
        } else if (m.type == "MixMaterial") {
            e.type = AKB_MAT_MIX;
            const Value *fr = m.find("fraction"), *a = m.find("first"), *b = m.find("second");
            if (!fr || !a || !b) bad(m, "fields fraction, first and second are required");
            float ff[3];
            e.fraction_tex = texture_rgb(m, "fraction", *fr, ff);
            e.fraction = ff[0];
            e.child_a = material_index(material_node(m, "first", *a));
            e.child_b = material_index(material_node(m, "second", *b));
        } else if (m.type == "DielectricMaterial") {
            e.type = AKB_MAT_DIELECTRIC;
            e.color[0] = e.color[1] = e.color[2] = 1.0f;
            e.ior = 1.5f;
            if (const Value *c = m.find("color"))
                if (texture_rgb(m, "color", *c, e.color) >= 0) bad(m, "field color of a DielectricMaterial must be a constant");
            if (const Value *i = m.find("ior")) e.ior = (float)num(m, "ior", *i);
        } else {
            bad(m, "not a material node");
        }
        sf->materials[id] = e;
        return id;
    }

    // AkariMesh::object_field + compile, core/nodes/mesh.cpp:30-87
    void mesh_from(const Object &o) {
        const Value *path = o.find("path");
        if (!path || !path->is_string()) bad(o, "field path (string) is required");
        auto md = std::make_unique<MeshData>();
        std::string file = join_path(dir_of(o.file), path->str);
        {
            FILE *probe = fopen(file.c_str(), "rb");
            if (probe)
                fclose(probe);
            else
                file = path->str; // the reference opens it relative to the working directory
        }
        read_mesh_file(file, md.get());
        if (const Value *mats = o.find("materials")) {
            if (!mats->is_array()) bad(o, "field materials must be an array");
            for (const Value &m : mats->arr) md->material_remap.push_back(material_index(material_node(o, "materials", m)));
        }
        akb_mesh d;
        memset(&d, 0, sizeof d);
        d.n_vertices = md->vertices.size() / 3, d.n_triangles = md->indices.size() / 3;
        d.vertices = md->vertices.data(), d.normals = md->normals.data(), d.texcoords = md->texcoords.data();
        d.indices = md->indices.data(), d.material_indices = md->material_indices.data();
        d.materials = md->material_remap.data(), d.n_mesh_materials = (int32_t)md->material_remap.size();
        sf->mesh_descs.push_back(d);
        sf->meshes.push_back(std::move(md));
    }
};

bool boolean(const Object &o, const char *field, const Value &v) {
    if (v.kind != Value::Bool) bad(o, std::string("field ") + field + " must be true or false");
    return v.b;
}

// PathIntegratorNode / AOIntegratorNode, core/nodes/integrator.cpp:26-84 (defaults included)
void integrator_from(const Object &o, SceneFile *sf) {
    akb_render_params &p = sf->params;
    memset(&p, 0, sizeof p);
    p.world_size = 1, p.tile_size = 32;
    p.sampler = AKB_SAMPLER_LCG; // core/nodes/scene.cpp:126
    if (o.type == "Path") {
        p.integrator = AKB_INTEGRATOR_PATH;
        p.spp = 16, p.max_depth = 5, p.ray_clamp = 10.0f, p.mode = AKB_MODE_WAVEFRONT;
        sf->tile_size = 256;
        for (auto &f : o.fields) {
            const Value &v = f.second;
            if (f.first == "spp") p.spp = (int)num(o, "spp", v);
            else if (f.first == "max_depth") p.max_depth = (int)num(o, "max_depth", v);
            else if (f.first == "tile_size") sf->tile_size = (int)num(o, "tile_size", v);
            else if (f.first == "ray_clamp") p.ray_clamp = (float)num(o, "ray_clamp", v);
            else if (f.first == "wavefront") p.mode = boolean(o, "wavefront", v) ? AKB_MODE_WAVEFRONT : AKB_MODE_MEGAKERNEL;
            else if (f.first == "megakernel") p.mode = boolean(o, "megakernel", v) ? AKB_MODE_MEGAKERNEL : AKB_MODE_WAVEFRONT;
            // extensions
            else if (f.first == "rr_start_depth") p.rr_start_depth = (int)num(o, "rr_start_depth", v);
            else if (f.first == "sort_materials") p.sort_materials = boolean(o, "sort_materials", v) ? 1 : 0;
            else if (f.first == "sampler") {
                if (!v.is_string() || (v.str != "lcg" && v.str != "pcg")) bad(o, "field sampler must be \"lcg\" or \"pcg\"");
                p.sampler = v.str == "pcg" ? AKB_SAMPLER_PCG : AKB_SAMPLER_LCG;
            }
        }
    } else if (o.type == "AO") {
        p.integrator = AKB_INTEGRATOR_AO;
        p.spp = 16, p.ao_occlude = INFINITY;
        sf->tile_size = 16;
        for (auto &f : o.fields) {
            if (f.first == "spp") p.spp = (int)num(o, "spp", f.second);
            else if (f.first == "occlude") p.ao_occlude = (float)num(o, "occlude", f.second);
        }
    } else {
        bad(o, "is not an integrator");
    }
}

} // namespace

void read_mesh_file(const std::string &path, MeshData *out) {
    static const char magic[] = "AKARI_BINARY_MESH"; // 17 bytes, no NUL on disk
    FILE *f = fopen(path.c_str(), "rb");
    if (!f) throw std::runtime_error("error loading " + path + ": cannot open");
    auto fail = [&](const char *why) {
        fclose(f);
        throw std::runtime_error("error loading " + path + ": " + why);
    };
    char m[17];
    uint64_t nv = 0, nt = 0;
    if (fread(m, 1, 17, f) != 17 || memcmp(m, magic, 17) != 0) fail("not an AKARI_BINARY_MESH file");
    if (fread(&nv, 8, 1, f) != 1 || fread(&nt, 8, 1, f) != 1) fail("truncated header");
    if (nv > (1ull << 33) || nt > (1ull << 33)) fail("implausible counts");
    out->vertices.resize(3 * nv), out->normals.resize(9 * nt), out->texcoords.resize(6 * nt);
    out->indices.resize(3 * nt), out->material_indices.resize(nt);
    auto rd = [&](void *p, size_t bytes) {
        if (bytes && fread(p, 1, bytes, f) != bytes) fail("truncated data");
    };
    rd(out->vertices.data(), 12 * nv);
    rd(out->normals.data(), 36 * nt);
    rd(out->texcoords.data(), 24 * nt);
    rd(out->indices.data(), 12 * nt);
    rd(out->material_indices.data(), 4 * nt);
    if (fread(m, 1, 17, f) != 17 || memcmp(m, magic, 17) != 0) fail("trailing magic missing");
    fclose(f);
}

std::unique_ptr<SceneFile> load_scene_file(const std::string &path) {
    auto mod = parse_file(path, "this", known_type);
    auto it = mod->exports.find("scene"); // cmd/akari.cpp:84-90
    if (it == mod->exports.end()) throw std::runtime_error(path + ": variable 'scene' not found");
    if (!it->second.is_object() || it->second.obj->type != "Scene") throw std::runtime_error(path + ": 'scene' is not a Scene node");
    const Object &scene = *it->second.obj;
    auto sf = std::make_unique<SceneFile>();
    Flattener fl{sf.get()};
    // SceneNode::object_field, core/nodes/scene.cpp:96-112
    const Value *cam = scene.find("camera"), *integ = scene.find("integrator"), *shapes = scene.find("shapes");
    if (!cam || !cam->is_object() || cam->obj->type != "PerspectiveCamera") bad(scene, "field camera must be a PerspectiveCamera");
    if (!integ || !integ->is_object()) bad(scene, "field integrator must be a Path or AO node");
    integrator_from(*integ->obj, sf.get());
    if (const Value *out = scene.find("output")) {
        if (!out->is_string()) bad(scene, "field output must be a string");
        sf->output = out->str;
    }
    if (shapes) {
        if (!shapes->is_array()) bad(scene, "field shapes must be an array");
        for (const Value &s : shapes->arr) {
            if (!s.is_object() || s.obj->type != "AkariMesh") bad(scene, "shapes must be AkariMesh nodes");
            fl.mesh_from(*s.obj);
        }
    }
    memset(&sf->desc, 0, sizeof sf->desc);
    sf->desc.abi_version = AKB_ABI_VERSION;
    sf->desc.n_meshes = (uint32_t)sf->mesh_descs.size(), sf->desc.meshes = sf->mesh_descs.data();
    sf->desc.n_materials = (uint32_t)sf->materials.size(), sf->desc.materials = sf->materials.data();
    sf->desc.camera = camera_from(*cam->obj);
    sf->desc.n_textures = (uint32_t)sf->textures.size(), sf->desc.textures = sf->textures.data();
    return sf;
}

// ---------------------------------------------------------------------------------------------------------------------
// PNG: 8-bit RGB, filter 0 on every row, zlib stream of stored (uncompressed) deflate blocks.
namespace {
uint32_t crc_table[256];
void crc_init() {
    if (crc_table[1]) return;
    for (uint32_t n = 0; n < 256; n++) {
        uint32_t c = n;
        for (int k = 0; k < 8; k++) c = (c & 1) ? 0xEDB88320u ^ (c >> 1) : c >> 1;
        crc_table[n] = c;
    }
}
uint32_t crc32(uint32_t crc, const uint8_t *p, size_t n) {
    crc = ~crc;
    for (size_t i = 0; i < n; i++) crc = crc_table[(crc ^ p[i]) & 0xFF] ^ (crc >> 8);
    return ~crc;
}
void be32(std::vector<uint8_t> &v, uint32_t x) {
    for (int s = 24; s >= 0; s -= 8) v.push_back((uint8_t)(x >> s));
}
void chunk(FILE *f, const char *tag, const std::vector<uint8_t> &data) {
    std::vector<uint8_t> head;
    be32(head, (uint32_t)data.size());
    fwrite(head.data(), 1, 4, f);
    std::vector<uint8_t> body(tag, tag + 4);
    body.insert(body.end(), data.begin(), data.end());
    fwrite(body.data(), 1, body.size(), f);
    std::vector<uint8_t> tail;
    be32(tail, crc32(0, body.data(), body.size()));
    fwrite(tail.data(), 1, 4, f);
}
} // namespace

void write_png_rgb8(const std::string &path, const uint8_t *rgb, int width, int height) {
    crc_init();
    FILE *f = fopen(path.c_str(), "wb");
    if (!f) throw std::runtime_error("cannot write " + path);
    static const uint8_t sig[8] = {0x89, 'P', 'N', 'G', 0x0D, 0x0A, 0x1A, 0x0A};
    fwrite(sig, 1, 8, f);
    std::vector<uint8_t> ihdr;
    be32(ihdr, (uint32_t)width), be32(ihdr, (uint32_t)height);
    ihdr.insert(ihdr.end(), {8, 2, 0, 0, 0}); // 8 bits, truecolour, deflate, adaptive filtering, no interlace
    chunk(f, "IHDR", ihdr);
    std::vector<uint8_t> raw;
    raw.reserve((size_t)height * (3 * (size_t)width + 1));
    for (int y = 0; y < height; y++) {
        raw.push_back(0);
        raw.insert(raw.end(), rgb + (size_t)y * width * 3, rgb + (size_t)(y + 1) * width * 3);
    }
    std::vector<uint8_t> z{0x78, 0x01};
    uint32_t a = 1, b = 0;
    size_t off = 0;
    do {
        const size_t n = std::min<size_t>(65535, raw.size() - off);
        z.push_back(off + n >= raw.size() ? 1 : 0);
        z.push_back((uint8_t)(n & 0xFF)), z.push_back((uint8_t)(n >> 8));
        z.push_back((uint8_t)(~n & 0xFF)), z.push_back((uint8_t)((~n >> 8) & 0xFF));
        z.insert(z.end(), raw.begin() + off, raw.begin() + off + n);
        for (size_t i = off; i < off + n; i++) {
            a = (a + raw[i]) % 65521u;
            b = (b + a) % 65521u;
        }
        off += n;
    } while (off < raw.size());
    be32(z, (b << 16) | a);
    chunk(f, "IDAT", z);
    chunk(f, "IEND", {});
    fclose(f);
}

void write_pfm_rgb(const std::string &path, const float *rgbw, int width, int height) {
    FILE *f = fopen(path.c_str(), "wb");
    if (!f) throw std::runtime_error("cannot write " + path);
    fprintf(f, "PF\n%d %d\n-1.0\n", width, height);
    std::vector<float> row(3 * (size_t)width);
    for (int y = height - 1; y >= 0; y--) { // PFM rows run bottom to top
        for (int x = 0; x < width; x++) {
            const float *p = rgbw + 4 * ((size_t)y * width + x);
            const float w = p[3] != 0.0f ? p[3] : 1.0f; // Film::write_image: radiance / weight where weight != 0
            row[3 * x] = p[0] / w, row[3 * x + 1] = p[1] / w, row[3 * x + 2] = p[2] / w;
        }
        fwrite(row.data(), sizeof(float), row.size(), f);
    }
    fclose(f);
}

} // namespace akb_host
