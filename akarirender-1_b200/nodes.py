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


# ---- materials (core/nodes/material.cpp:89-170) ------------------------------------------------------------
class MaterialNode:
    type = None

    def entry(self, index_of):
        raise NotImplementedError


def _rgb(v):
    if isinstance(v, (int, float)):
        return (float(v),) * 3
    return tuple(float(x) for x in v)


class DiffuseMaterial(MaterialNode):
    def __init__(self, color):
        self.color = _rgb(color)

    def entry(self, index_of):
        return abi.Material(abi.MAT_DIFFUSE, (C.c_float * 3)(*self.color), 0, 0, -1, -1, 0, 0)


class GlossyMaterial(MaterialNode):
    def __init__(self, color, roughness):
        self.color, self.roughness = _rgb(color), float(roughness)

    def entry(self, index_of):
        return abi.Material(abi.MAT_GLOSSY, (C.c_float * 3)(*self.color), self.roughness, 0, -1, -1, 0, 0)


class EmissiveMaterial(MaterialNode):
    def __init__(self, color, double_sided=False):
        # the reference node never sets double_sided (core/nodes/material.cpp:152-155)
        self.color, self.double_sided = _rgb(color), bool(double_sided)

    def entry(self, index_of):
        return abi.Material(abi.MAT_EMISSIVE, (C.c_float * 3)(*self.color), 0, 0, -1, -1, int(self.double_sided), 0)


class MixMaterial(MaterialNode):
    def __init__(self, fraction, first, second):
        self.fraction, self.first, self.second = float(fraction), first, second

    def entry(self, index_of):
        return abi.Material(abi.MAT_MIX, (C.c_float * 3)(0, 0, 0), 0, self.fraction, index_of(self.first),
                            index_of(self.second), 0, 0)


class DielectricMaterial(MaterialNode):
    """Extension: smooth dielectric; not present in the reference."""

    def __init__(self, color=(1, 1, 1), ior=1.5):
        self.color, self.ior = _rgb(color), float(ior)

    def entry(self, index_of):
        return abi.Material(abi.MAT_DIELECTRIC, (C.c_float * 3)(*self.color), 0, 0, -1, -1, 0, self.ior)


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
    lines, names = [], {}

    def material(m):
        if id(m) in names:
            return names[id(m)]
        if isinstance(m, MixMaterial):
            a, b = material(m.first), material(m.second)
            body = "MixMaterial { fraction: %s, first: $%s, second: $%s }" % (sdl_number(m.fraction), a, b)
        elif isinstance(m, DiffuseMaterial):
            body = "DiffuseMaterial { color: %s }" % _sdl_list(m.color)
        elif isinstance(m, GlossyMaterial):
            body = "GlossyMaterial { color: %s, roughness: %s }" % (_sdl_list(m.color), sdl_number(m.roughness))
        elif isinstance(m, EmissiveMaterial):
            body = "EmissiveMaterial { color: %s }" % _sdl_list(m.color)
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
            mats_entry = table[i].entry(index_of)
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
        keep += [meshes, mats]
        return d, keep
