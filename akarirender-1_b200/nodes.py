"""Host-side mirror of the reference's scene-graph node layer for the Path hot path.

Same node names, field names, units and defaults as the reference's registered nodes
(core/nodes/{scene,camera,mesh,material,integrator}.cpp), flattened into the
`akb_scene_desc` of include/akari_b200.h the way SceneNode::compile / MeshNode::compile do
(core/nodes/scene.cpp:43-95, core/nodes/mesh.cpp:47-61). Used by tests and bench.py; the C++ front-end
(host/) does the same for `.akari` files.
"""
import ctypes as C
import math
import struct

import numpy as np

from . import abi

F32 = np.float32


# ---- textures + materials (core/nodes/material.cpp:25-170) ---------------------------------------------------
class ImageTexture:
    """ImageTextureNode (core/nodes/material.cpp:46-71): texels as the reference's reader stores them
    (core/image.cpp:105-123: byte / 255.0f as float32, alpha 1). `rgba` is (H, W, 4) float32, top row first."""

    def __init__(self, path=None, rgba=None):
        self.path = path
        if rgba is None:
            rgb8 = read_png_rgb8(path)
            rgba = np.ones(rgb8.shape[:2] + (4,), F32)
            rgba[..., :3] = rgb8.astype(F32) / F32(255.0)
        self.rgba = np.ascontiguousarray(rgba, F32)
        assert self.rgba.ndim == 3 and self.rgba.shape[2] == 4

    def rgb8(self):
        return np.clip(np.rint(self.rgba[..., :3] * 255.0), 0, 255).astype(np.uint8)


def read_png_rgb8(path):
    """8-bit non-interlaced PNG -> (H, W, 3) uint8 (grey and alpha variants are expanded / dropped like stbi_load(..., 3))."""
    import zlib
    b = open(path, "rb").read()
    assert b[:8] == b"\x89PNG\r\n\x1a\n", "not a PNG file"
    o, idat, hdr = 8, b"", None
    while o < len(b):
        n, tag = struct.unpack(">I4s", b[o:o + 8])
        if tag == b"IHDR":
            hdr = struct.unpack(">IIBBBBB", b[o + 8:o + 8 + n])
        elif tag == b"IDAT":
            idat += b[o + 8:o + 8 + n]
        o += 12 + n
    w, h, depth, ctype, _, _, lace = hdr
    assert depth == 8 and lace == 0 and ctype in (0, 2, 4, 6), "unsupported PNG variant"
    ch = {0: 1, 2: 3, 4: 2, 6: 4}[ctype]
    raw = np.frombuffer(zlib.decompress(idat), np.uint8).reshape(h, w * ch + 1)
    out = np.zeros((h, w * ch), np.int32)
    for y in range(h):
        f, line = raw[y, 0], raw[y, 1:].astype(np.int32)
        prev = out[y - 1] if y else np.zeros(w * ch, np.int32)
        if f == 0:
            cur = line
        elif f == 2:
            cur = (line + prev) & 255
        else:
            cur = np.zeros(w * ch, np.int32)
            for i in range(w * ch):
                a = cur[i - ch] if i >= ch else 0
                c = prev[i - ch] if i >= ch else 0
                if f == 1:
                    pr = a
                elif f == 3:
                    pr = (a + prev[i]) >> 1
                else:
                    pa, pb, pc = abs(prev[i] - c), abs(a - c), abs(a + prev[i] - 2 * c)
                    pr = a if pa <= pb and pa <= pc else (prev[i] if pb <= pc else c)
                cur[i] = (line[i] + pr) & 255
        out[y] = cur
    px = out.reshape(h, w, ch).astype(np.uint8)
    if ch <= 2:
        return np.repeat(px[..., :1], 3, axis=2)
    return px[..., :3].copy()


def write_png_rgb8(path, rgb8):
    import zlib
    h, w = rgb8.shape[:2]
    raw = b"".join(b"\x00" + rgb8[y].tobytes() for y in range(h))

    def chunk(tag, body):
        return struct.pack(">I", len(body)) + tag + body + struct.pack(">I", zlib.crc32(tag + body))
    with open(path, "wb") as f:
        f.write(b"\x89PNG\r\n\x1a\n" + chunk(b"IHDR", struct.pack(">IIBBBBB", w, h, 8, 2, 0, 0, 0))
                + chunk(b"IDAT", zlib.compress(raw, 6)) + chunk(b"IEND", b""))


class MaterialNode:
    type = None

    def entry(self, index_of, tex_of):
        raise NotImplementedError


def _rgb(v):
    """A texture value as the scene language has it: [r,g,b] | number | ImageTexture (core/nodes/material.cpp:73-88)."""
    if isinstance(v, ImageTexture):
        return v
    if isinstance(v, (int, float)):
        return (float(v),) * 3
    return tuple(float(x) for x in v)


def _const(v):
    return (0.0, 0.0, 0.0) if isinstance(v, ImageTexture) else v


def _mat(type_, color=(0, 0, 0), roughness=0.0, fraction=0.0, a=-1, b=-1, double_sided=0, ior=0.0, tex=(-1, -1, -1)):
    return abi.Material(type_, (C.c_float * 3)(*color), roughness, fraction, a, b, double_sided, ior, *tex)


class DiffuseMaterial(MaterialNode):
    def __init__(self, color):
        self.color = _rgb(color)

    def entry(self, index_of, tex_of):
        return _mat(abi.MAT_DIFFUSE, _const(self.color), tex=(tex_of(self.color), -1, -1))


class GlossyMaterial(MaterialNode):
    def __init__(self, color, roughness):
        self.color = _rgb(color)
        self.roughness = roughness if isinstance(roughness, ImageTexture) else float(roughness)

    def entry(self, index_of, tex_of):
        r = 0.0 if isinstance(self.roughness, ImageTexture) else self.roughness
        return _mat(abi.MAT_GLOSSY, _const(self.color), roughness=r, tex=(tex_of(self.color), tex_of(self.roughness), -1))


class EmissiveMaterial(MaterialNode):
    def __init__(self, color, double_sided=False):
        # the reference node never sets double_sided (core/nodes/material.cpp:152-155)
        self.color, self.double_sided = _rgb(color), bool(double_sided)

    def entry(self, index_of, tex_of):
        return _mat(abi.MAT_EMISSIVE, _const(self.color), double_sided=int(self.double_sided), tex=(tex_of(self.color), -1, -1))


class MixMaterial(MaterialNode):
    def __init__(self, fraction, first, second):
        self.fraction = fraction if isinstance(fraction, ImageTexture) else float(fraction)
        self.first, self.second = first, second

    def entry(self, index_of, tex_of):
        f = 0.0 if isinstance(self.fraction, ImageTexture) else self.fraction
        return _mat(abi.MAT_MIX, fraction=f, a=index_of(self.first), b=index_of(self.second), tex=(-1, -1, tex_of(self.fraction)))


class DielectricMaterial(MaterialNode):
    """Extension: smooth dielectric; not present in the reference."""

    def __init__(self, color=(1, 1, 1), ior=1.5):
        self.color, self.ior = tuple(float(x) for x in _rgb(color)), float(ior)

    def entry(self, index_of, tex_of):
        return _mat(abi.MAT_DIELECTRIC, self.color, ior=self.ior)


# ---- camera (core/nodes/camera.cpp:26-51) -------------------------------------------------------------------
def _matmul4(a, b):
    """Matrix::operator* (common/math.h:97-105): m(i,j) = dot(row i, col j), left-to-right float adds."""
    m = np.zeros((4, 4), F32)
    for i in range(4):
        for j in range(4):
            s = F32(a[i, 0] * b[0, j])
            for k in range(1, 4):
                s = F32(s + F32(a[i, k] * b[k, j]))
            m[i, j] = s
    return m


PI_F = F32(3.1415926535897932384)
_libm_handle = None


def _libm():
    global _libm_handle
    if _libm_handle is None:
        L = C.CDLL("libm.so.6")
        for f in (L.sinf, L.cosf):
            f.restype, f.argtypes = C.c_float, [C.c_float]
        _libm_handle = L
    return _libm_handle


class PerspectiveCamera:
    def __init__(self, fov=80.0, position=(0, 0, 0), rotation=(0, 0, 0), resolution=(512, 512)):
        self.fov = float(fov)  # degrees, as in the scene file
        self.position = tuple(float(x) for x in position)
        self.rotation = tuple(float(x) for x in rotation)  # degrees: (yaw, pitch, roll)
        self.resolution = (int(resolution[0]), int(resolution[1]))

    def c2w(self):
        # radians(float3) in float (common/math.h:351-354, Pi() is a float literal); sin/cos are glibc's sinf/cosf
        rx, ry, rz = (F32(F32(F32(r) * PI_F) / F32(180.0)) for r in self.rotation)

        def rot(axis, th):
            s, c = F32(_libm().sinf(float(th))), F32(_libm().cosf(float(th)))
            m = np.eye(4, dtype=F32)
            if axis == "x":
                m[1, 1], m[1, 2], m[2, 1], m[2, 2] = c, -s, s, c
            elif axis == "y":
                m[0, 0], m[0, 2], m[2, 0], m[2, 2] = c, s, -s, c
            else:
                m[0, 0], m[0, 1], m[1, 0], m[1, 1] = c, -s, s, c
            return m

        t = np.eye(4, dtype=F32)
        t[0, 3], t[1, 3], t[2, 3] = self.position
        m = rot("z", rz)
        m = _matmul4(rot("x", ry), m)
        m = _matmul4(rot("y", rx), m)
        m = _matmul4(t, m)
        return m

    def desc(self):
        cam = abi.Camera()
        cam.c2w[:] = [float(x) for x in self.c2w().reshape(-1)]
        # `fov = radians(value.get<double>())`: double arithmetic with the float-literal Pi, narrowed by the camera ctor
        cam.fov = float(F32(self.fov * float(PI_F) / 180.0))
        cam.width, cam.height = self.resolution
        cam.lens_radius = 0.0
        cam.focal_distance = 0.0
        return cam


# ---- mesh (core/nodes/mesh.cpp, core/mesh.cpp:27-85) --------------------------------------------------------
MESH_MAGIC = b"AKARI_BINARY_MESH"


class AkariMesh:
    def __init__(self, path=None, materials=(), vertices=None, normals=None, texcoords=None, indices=None,
                 material_indices=None):
        self.materials = list(materials)
        if path is not None:
            vertices, normals, texcoords, indices, material_indices = read_mesh(path)
        self.vertices = np.ascontiguousarray(vertices, F32).reshape(-1)
        self.indices = np.ascontiguousarray(indices, np.int32).reshape(-1)
        nt = self.indices.size // 3
        if normals is None:
            normals = flat_normals(self.vertices, self.indices)
        if texcoords is None:
            # importer default (cmd/akari-import.cpp:85-87)
            texcoords = np.tile(np.array([0, 1, 1, 0, 1, 1], F32), nt)
        self.normals = np.ascontiguousarray(normals, F32).reshape(-1)
        self.texcoords = np.ascontiguousarray(texcoords, F32).reshape(-1)
        self.material_indices = np.ascontiguousarray(material_indices, np.int32).reshape(-1)
        assert self.normals.size == 9 * nt and self.texcoords.size == 6 * nt and self.material_indices.size == nt

    @property
    def n_triangles(self):
        return self.indices.size // 3


def flat_normals(vertices, indices):
    v = vertices.reshape(-1, 3)[indices.reshape(-1, 3)]
    n = np.cross(v[:, 1] - v[:, 0], v[:, 2] - v[:, 0])
    n = n / np.maximum(np.linalg.norm(n, axis=1, keepdims=True), 1e-30)
    return np.repeat(n.astype(F32), 3, axis=0).reshape(-1)


def read_mesh(path):
    with open(path, "rb") as f:
        b = f.read()
    assert b[:17] == MESH_MAGIC, "not an AKARI_BINARY_MESH file"
    nv, nt = struct.unpack_from("<QQ", b, 17)
    o = 33
    out = []
    for cnt, dt in ((3 * nv, F32), (9 * nt, F32), (6 * nt, F32), (3 * nt, np.int32), (nt, np.int32)):
        out.append(np.frombuffer(b, dt, cnt, o).copy())
        o += 4 * cnt
    assert b[o:o + 17] == MESH_MAGIC
    return out


def write_mesh(path, mesh):
    with open(path, "wb") as f:
        f.write(MESH_MAGIC)
        f.write(struct.pack("<QQ", mesh.vertices.size // 3, mesh.n_triangles))
        for a in (mesh.vertices, mesh.normals, mesh.texcoords, mesh.indices, mesh.material_indices):
            f.write(a.tobytes())
        f.write(MESH_MAGIC)


# ---- integrators (core/nodes/integrator.cpp:26-90) ----------------------------------------------------------
class Path:
    def __init__(self, spp=16, max_depth=5, tile_size=256, ray_clamp=10.0, wavefront=True, megakernel=False,
                 sampler="lcg", rr_start_depth=0, sort_materials=False):
        self.spp, self.max_depth, self.tile_size, self.ray_clamp = int(spp), int(max_depth), int(tile_size), float(
            ray_clamp)
        self.wavefront = bool(wavefront) and not megakernel
        self.sampler, self.rr_start_depth, self.sort_materials = sampler, int(rr_start_depth), bool(sort_materials)

    def params(self, rank=0, world_size=1, shard_tile=32, split=abi.SPLIT_TILE):
        p = abi.RenderParams()
        p.split = split
        p.integrator = abi.INTEGRATOR_PATH
        p.spp, p.max_depth, p.ray_clamp = self.spp, self.max_depth, self.ray_clamp
        p.sampler = abi.SAMPLER_PCG if self.sampler == "pcg" else abi.SAMPLER_LCG
        p.rr_start_depth = self.rr_start_depth
        p.mode = abi.MODE_WAVEFRONT if self.wavefront else abi.MODE_MEGAKERNEL
        p.sort_materials = int(self.sort_materials)
        p.ao_occlude = 0.0
        p.rank, p.world_size, p.tile_size = rank, world_size, shard_tile
        return p


class AO:
    def __init__(self, spp=16, occlude=float("inf")):
        self.spp, self.occlude = int(spp), float(occlude)

    def params(self, rank=0, world_size=1, shard_tile=32, split=abi.SPLIT_TILE):
        p = abi.RenderParams()
        p.split = split
        p.integrator = abi.INTEGRATOR_AO
        p.spp, p.max_depth, p.ray_clamp = self.spp, 0, 0.0
        p.sampler, p.rr_start_depth, p.mode, p.sort_materials = abi.SAMPLER_LCG, 0, abi.MODE_MEGAKERNEL, 0
        p.ao_occlude = self.occlude
        p.rank, p.world_size, p.tile_size = rank, world_size, shard_tile
        return p


# ---- `.akari` text (core/parser.cpp) --------------------------------------------------
def sdl_number(v):
    """Shortest fixed-point literal (the SDL has no exponent syntax) that the reference's number reader
    (core/parser.cpp:338-363: n + frac / 10^k in double, then narrowed) maps back to float32(v)."""
    f = F32(v)
    if not np.isfinite(f):
        raise ValueError("the scene language cannot express %r" % (v,))
    for prec in range(1, 60):
        t = np.format_float_positional(f, precision=prec, unique=False, fractional=False, trim="-")
        neg, body = t.startswith("-"), t.lstrip("-")
        ip, _, fp = body.partition(".")
        if len(ip.lstrip("0")) + len(fp.lstrip("0")) > 18:
            break
        val = float(int(ip)) + (float(int(fp)) / float(10 ** len(fp)) if fp else 0.0)
        if F32(-val if neg else val) == f:
            return t
    raise ValueError("no exact literal for %r" % (v,))


def _sdl_list(vals):
    return "[" + ", ".join(sdl_number(x) for x in vals) + "]"


def write_akari(scene, directory, name="scene.akari"):
    """Writes `scene` as a reference-format scene file + one .mesh per shape into `directory`. Returns the path."""
    import os
    os.makedirs(directory, exist_ok=True)
    lines, names, tex_files = [], {}, {}

    def tex(v):
        """A texture slot's literal: [r,g,b] / number, or the path of a PNG written next to the scene file."""
        if isinstance(v, ImageTexture):
            if id(v) not in tex_files:
                tex_files[id(v)] = "tex%d.png" % len(tex_files)
                write_png_rgb8(os.path.join(directory, tex_files[id(v)]), v.rgb8())
            return '"%s"' % tex_files[id(v)]
        return _sdl_list(v) if isinstance(v, tuple) else sdl_number(v)

    def material(m):
        if id(m) in names:
            return names[id(m)]
        if isinstance(m, MixMaterial):
            a, b = material(m.first), material(m.second)
            body = "MixMaterial { fraction: %s, first: $%s, second: $%s }" % (tex(m.fraction), a, b)
        elif isinstance(m, DiffuseMaterial):
            body = "DiffuseMaterial { color: %s }" % tex(m.color)
        elif isinstance(m, GlossyMaterial):
            body = "GlossyMaterial { color: %s, roughness: %s }" % (tex(m.color), tex(m.roughness))
        elif isinstance(m, EmissiveMaterial):
            body = "EmissiveMaterial { color: %s }" % tex(m.color)
        elif isinstance(m, DielectricMaterial):
            body = "DielectricMaterial { color: %s, ior: %s }" % (_sdl_list(m.color), sdl_number(m.ior))
        else:
            raise TypeError(m)
        names[id(m)] = "mat%d" % len(names)
        lines.append("export %s = %s" % (names[id(m)], body))
        return names[id(m)]

    shape_names = []
    for k, sh in enumerate(scene.shapes):
        write_mesh(os.path.join(directory, "shape%d.mesh" % k), sh)
        mats = ", ".join("$" + material(m) for m in sh.materials)
        lines.append("export shape%d = AkariMesh {\n  path: \"shape%d.mesh\",\n  materials: [%s]\n}" % (k, k, mats))
        shape_names.append("$shape%d" % k)
    cam = scene.camera
    lines.append("export camera = PerspectiveCamera {\n  fov: %s,\n  position: %s,\n  rotation: %s,\n  resolution: [%d, %d]\n}"
                 % (sdl_number(cam.fov), _sdl_list(cam.position), _sdl_list(cam.rotation), cam.resolution[0],
                    cam.resolution[1]))
    it = scene.integrator
    if isinstance(it, AO):
        integ = "AO { spp: %d%s }" % (it.spp, "" if math.isinf(it.occlude) else ", occlude: " + sdl_number(it.occlude))
    else:
        integ = ("Path { spp: %d, max_depth: %d, tile_size: %d, ray_clamp: %s, wavefront: %s"
                 % (it.spp, it.max_depth, it.tile_size, sdl_number(it.ray_clamp), "true" if it.wavefront else "false"))
        if it.rr_start_depth:
            integ += ", rr_start_depth: %d" % it.rr_start_depth
        if it.sampler == "pcg":
            integ += ", sampler: \"pcg\""
        if it.sort_materials:
            integ += ", sort_materials: true"
        integ += " }"
    lines.append("// written by akarirender-1_b200.nodes.write_akari\nexport scene = Scene {\n  camera: $camera,\n  integrator: %s,\n"
                 "  output: \"%s\",\n  shapes: [%s]\n}" % (integ, scene.output, ", ".join(shape_names)))
    path = os.path.join(directory, name)
    with open(path, "w") as f:
        f.write("\n".join(lines) + "\n")
    return path


class SceneFile:
    """A scene parsed by the C++ front-end (akb_scene_file_*): the `akari scene.akari` entry of the reference
    (cmd/akari.cpp:78-92). Quacks like `Scene` for `Backend` (flatten / camera.resolution / integrator.params)."""

    def __init__(self, path):
        self.lib = abi.load()
        h = C.c_void_p()
        if self.lib.akb_scene_file_load(str(path).encode(), C.byref(h)) != abi.OK:
            raise RuntimeError(self.lib.akb_last_error().decode(errors="replace"))
        self.h = h
        self.desc = self.lib.akb_scene_file_desc(h).contents
        self.output = self.lib.akb_scene_file_output(h).decode()
        self.camera = self
        self.integrator = self
        self.resolution = (self.desc.camera.width, self.desc.camera.height)

    def params(self, rank=0, world_size=1, shard_tile=32, split=abi.SPLIT_TILE):
        p = abi.RenderParams()
        if self.lib.akb_scene_file_params(self.h, C.byref(p)) != abi.OK:
            raise RuntimeError(self.lib.akb_last_error().decode(errors="replace"))
        p.rank, p.world_size, p.tile_size, p.split = rank, world_size, shard_tile, split
        return p

    @property
    def spp(self):
        return self.params().spp

    def flatten(self):
        return self.desc, [self]

    def close(self):
        if getattr(self, "h", None):
            self.lib.akb_scene_file_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


# ---- scene (core/nodes/scene.cpp:96-165) --------------------------------------------------------------------
class Scene:
    def __init__(self, camera, integrator, shapes, output="out.png"):
        self.camera, self.integrator, self.shapes, self.output = camera, integrator, list(shapes), output

    def flatten(self):
        """-> (akb_scene_desc, keepalive). Materials are deduplicated by identity into one scene table."""
        table, ids = [], {}

        def index_of(node):
            if id(node) not in ids:
                ids[id(node)] = len(table)
                table.append(node)
                if isinstance(node, MixMaterial):
                    index_of(node.first), index_of(node.second)
            return ids[id(node)]

        textures, tex_ids = [], {}

        def tex_of(v):
            if not isinstance(v, ImageTexture):
                return -1
            if id(v) not in tex_ids:
                tex_ids[id(v)] = len(textures)
                textures.append(v)
            return tex_ids[id(v)]

        keep = []
        meshes = (abi.Mesh * len(self.shapes))()
        fp, ip = C.POINTER(C.c_float), C.POINTER(C.c_int32)
        for k, sh in enumerate(self.shapes):
            remap = np.array([index_of(m) for m in sh.materials], np.int32)
            keep += [sh.vertices, sh.normals, sh.texcoords, sh.indices, sh.material_indices, remap]
            m = meshes[k]
            m.n_vertices, m.n_triangles = sh.vertices.size // 3, sh.n_triangles
            m.vertices, m.normals, m.texcoords = (a.ctypes.data_as(fp) for a in (sh.vertices, sh.normals, sh.texcoords))
            m.indices, m.material_indices = sh.indices.ctypes.data_as(ip), sh.material_indices.ctypes.data_as(ip)
            m.materials, m.n_mesh_materials = remap.ctypes.data_as(ip), remap.size
        mats = (abi.Material * max(1, len(table)))()
        i = 0
        while i < len(table):  # table may grow while Mix children are indexed
            mats_entry = table[i].entry(index_of, tex_of)
            if len(table) > len(mats):
                new = (abi.Material * len(table))()
                for j in range(i):
                    new[j] = mats[j]
                mats = new
            mats[i] = mats_entry
            i += 1
        d = abi.SceneDesc()
        d.abi_version = abi.ABI_VERSION
        d.n_meshes, d.meshes = len(self.shapes), meshes
        d.n_materials, d.materials = len(table), mats
        d.camera = self.camera.desc()
        texs = (abi.Texture * max(1, len(textures)))()
        for i, t in enumerate(textures):
            texs[i].height, texs[i].width = t.rgba.shape[:2]
            texs[i].rgba = t.rgba.ctypes.data_as(fp)
        d.n_textures, d.textures = len(textures), texs
        keep += [meshes, mats, texs, textures]
        return d, keep
