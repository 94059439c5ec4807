"""Synthetic scenes of the shapes BASELINE.json names (C1..C5), built with the node mirror.

All geometry is generated from fixed seeds; there is no network and no dataset. `scale` shrinks the terrain grid
so the same scene family can be used at oracle-friendly sizes in tests.
"""
import os

import numpy as np

from .nodes import (AkariMesh, DielectricMaterial, DiffuseMaterial, EmissiveMaterial, GlossyMaterial, MixMaterial,
                    Path, PerspectiveCamera, Scene)

F32 = np.float32
_HERE = os.path.dirname(os.path.abspath(__file__))
CORNELL_FIXTURE = os.path.join(os.path.dirname(_HERE), "tests", "golden", "cornell_box_mesh.npz")


def cornell_box(resolution=(512, 512), spp=16, max_depth=5, **path_kw):
    """C1: the reference's shipped Cornell box (resources/data/cornell_box/{scene,cornell_box}.akari):
    36 triangles, 8 materials, camera fov 15 at (0,1,9). Geometry comes from the committed fixture
    tests/golden/cornell_box_mesh.npz (written by tests/golden/make_golden.py from the reference's .mesh)."""
    z = np.load(CORNELL_FIXTURE)
    white = (0.725, 0.71, 0.68)
    mats = [DiffuseMaterial((0.63, 0.065, 0.05)), DiffuseMaterial((0.14, 0.45, 0.091)), DiffuseMaterial(white),
            DiffuseMaterial(white), DiffuseMaterial(white), DiffuseMaterial(white), DiffuseMaterial(white),
            EmissiveMaterial((17, 12, 4))]
    mesh = AkariMesh(materials=mats, vertices=z["vertices"], normals=z["normals"], texcoords=z["texcoords"],
                     indices=z["indices"], material_indices=z["material_indices"])
    cam = PerspectiveCamera(fov=15, position=(0, 1, 9), rotation=(0, 0, 0), resolution=resolution)
    return Scene(cam, Path(spp=spp, max_depth=max_depth, ray_clamp=0.0, **path_kw), [mesh])


def terrain_arrays(n=708, extent=50.0, seed=1234, light=True, light_y=30.0, light_half=10.0):
    """M1: height field y = 3 sin(0.21x) cos(0.17z) + U(-0.05,0.05) on an n x n vertex grid over
    [-extent,extent]^2, two +y-facing triangles per cell, plus an optional 2-triangle light facing -y."""
    rng = np.random.default_rng(seed)
    xs = np.linspace(-extent, extent, n, dtype=np.float64)
    X, Z = np.meshgrid(xs, xs, indexing="ij")
    Y = 3.0 * np.sin(0.21 * X) * np.cos(0.17 * Z) + rng.uniform(-0.05, 0.05, X.shape)
    verts = np.stack([X, Y, Z], -1).reshape(-1, 3).astype(F32)
    i, j = np.meshgrid(np.arange(n - 1), np.arange(n - 1), indexing="ij")
    a = (i * n + j).reshape(-1)
    b, c, d = a + n, a + 1, a + n + 1  # b = (i+1,j), c = (i,j+1), d = (i+1,j+1)
    tris = np.concatenate([np.stack([a, c, b], -1), np.stack([b, c, d], -1)], 1).reshape(-1, 3).astype(np.int32)
    n_terrain = tris.shape[0]
    if light:
        h = light_half
        lv = np.array([[-h, light_y, -h], [h, light_y, -h], [-h, light_y, h], [h, light_y, h]], F32)
        base = verts.shape[0]
        verts = np.concatenate([verts, lv])
        tris = np.concatenate([tris, np.array([[base, base + 1, base + 2], [base + 1, base + 3, base + 2]], np.int32)])
    return verts, tris, n_terrain


def terrain_scene(n=708, resolution=(1920, 1080), spp=64, max_depth=8, materials="c3", seed=1234, light_radiance=40.0,
                  **path_kw):
    """C3: M1 with per-triangle materials 60% Diffuse / 30% Glossy(0.3) / 10% Dielectric(1.5) [extension];
    materials='c3ref' maps the dielectric 10% to Diffuse so the un-extended reference applies; 'diffuse' = all
    Diffuse (C5 base). Camera (0,20,70) pitch -15deg, fov 60."""
    verts, tris, n_terrain = terrain_arrays(n, seed=seed)
    rng = np.random.default_rng(seed)
    r = rng.random(n_terrain)
    diffuse = DiffuseMaterial((0.7, 0.6, 0.5))
    glossy = GlossyMaterial((0.8, 0.8, 0.8), 0.3)
    glass = DielectricMaterial((1, 1, 1), 1.5)
    emit = EmissiveMaterial((light_radiance,) * 3)
    if materials == "c3":
        mats = [diffuse, glossy, glass, emit]
        mi = np.where(r < 0.6, 0, np.where(r < 0.9, 1, 2)).astype(np.int32)
    elif materials == "c3ref":
        mats = [diffuse, glossy, DiffuseMaterial((0.7, 0.6, 0.5)), emit]
        mi = np.where(r < 0.6, 0, np.where(r < 0.9, 1, 2)).astype(np.int32)
    else:
        mats = [diffuse, glossy, glass, emit]
        mi = np.zeros(n_terrain, np.int32)
    mi = np.concatenate([mi, np.full(tris.shape[0] - n_terrain, 3, np.int32)])
    mesh = AkariMesh(materials=mats, vertices=verts, indices=tris, material_indices=mi)
    cam = PerspectiveCamera(fov=60, position=(0, 20, 70), rotation=(0, -15, 0), resolution=resolution)
    return Scene(cam, Path(spp=spp, max_depth=max_depth, ray_clamp=0.0, **path_kw), [mesh])


def instanced_scene(n=708, resolution=(3840, 2160), spp=256, max_depth=8, copies=(5, 2), pitch=110.0, **path_kw):
    """C4: copies[0] x copies[1] placed copies of M1 (translations baked: the reference has no instance transforms,
    so each copy is its own AkariMesh and the reference TLAS spans them)."""
    verts, tris, n_terrain = terrain_arrays(n)
    diffuse = DiffuseMaterial((0.7, 0.6, 0.5))
    glossy = GlossyMaterial((0.8, 0.8, 0.8), 0.3)
    emit = EmissiveMaterial((40.0,) * 3)
    rng = np.random.default_rng(1234)
    mi = np.where(rng.random(n_terrain) < 0.7, 0, 1).astype(np.int32)
    mi = np.concatenate([mi, np.full(tris.shape[0] - n_terrain, 2, np.int32)])
    shapes = []
    nx, nz = copies
    for ix in range(nx):
        for iz in range(nz):
            off = np.array([(ix - (nx - 1) / 2) * pitch, 0.0, (iz - (nz - 1) / 2) * pitch], F32)
            shapes.append(AkariMesh(materials=[diffuse, glossy, emit], vertices=verts + off, indices=tris,
                                    material_indices=mi))
    cam = PerspectiveCamera(fov=60, position=(0, 120, 330), rotation=(0, -25, 0), resolution=resolution)
    return Scene(cam, Path(spp=spp, max_depth=max_depth, ray_clamp=0.0, **path_kw), shapes)


def multi_mesh_scene(n=20, resolution=(64, 36), spp=4, max_depth=5, coincident=False, **path_kw):
    """C4 in miniature, for parity tests of the two-level wrapper (kernel/bvh-accelerator.h:651-678: geom_id = mesh
    index, prim_id = index inside the mesh, candidates of different meshes resolved by `localHit.t < record.t`):
    2 x 2 placed copies of a small M1 at a pitch of 90 units, so that neighbouring copies OVERLAP in 10-unit strips and
    rays cross from one mesh into the next; every mesh has its own material table in a different order (the per-mesh
    material remap). coincident=True puts the fourth mesh exactly on top of mesh 0 (the same triangles, so every hit on
    either is an exact-t tie between two meshes) plus one small triangle below the terrain.
    The scene stays at FOUR meshes on purpose: from five meshes on the reference's own TLAS loses meshes (measured with
    the reference build: 5x1, 3x2 and 5x2 placed copies are only ever hit on their last one or two meshes;
    BVHHandleConstructor, kernel/bvh-accelerator.h:643-648, stores the address of a constructor argument in every
    handle - a dangling pointer once the constructor returns), which is not behaviour to pin."""
    verts, tris, n_terrain = terrain_arrays(n)
    palette = [DiffuseMaterial((0.7, 0.6, 0.5)), GlossyMaterial((0.8, 0.8, 0.8), 0.3), DiffuseMaterial((0.2, 0.5, 0.8)),
               EmissiveMaterial((40.0,) * 3)]
    rng = np.random.default_rng(77)
    kind = rng.integers(0, 3, n_terrain)  # which palette entry a terrain triangle uses (3 = the light)
    kind = np.concatenate([kind, np.full(tris.shape[0] - n_terrain, 3)])
    shapes = []
    offs = [np.array([(ix - 0.5) * 90.0, 0.0, (iz - 0.5) * 90.0], F32) for ix in range(2) for iz in range(2)]
    if coincident:
        offs[3] = offs[0]
    for k, off in enumerate(offs):
        order = np.roll(np.arange(4), k)  # mesh k's table: palette[order[j]] at local index j
        local = np.argsort(order)         # palette entry -> local index
        v, t, mi = verts + off, tris, local[kind].astype(np.int32)
        if coincident and k == 3:
            extra = np.array([[0, -20, 0], [1, -20, 0], [0, -20, 1]], F32) + off
            t = np.concatenate([t, np.array([[v.shape[0], v.shape[0] + 1, v.shape[0] + 2]], np.int32)])
            v = np.concatenate([v, extra])
            mi = np.concatenate([mi, local[:1].astype(np.int32)])
        shapes.append(AkariMesh(materials=[palette[i] for i in order], vertices=v, indices=t, material_indices=mi))
    cam = PerspectiveCamera(fov=60, position=(0, 60, 150), rotation=(0, -25, 0), resolution=resolution)
    return Scene(cam, Path(spp=spp, max_depth=max_depth, ray_clamp=0.0, **path_kw), shapes)


def many_light_scene(n=708, resolution=(1920, 1080), spp=128, max_depth=16, n_quads=5000, rr_start_depth=3, seed=1234,
                     **path_kw):
    """C5: all-Diffuse M1 + n_quads emissive 0.5x0.5 quads on a 100x50 lattice at y = 25 + U(0,3), facing -y,
    radiance U(5,50) per quad; depth 16 with Russian roulette from depth 3 [extension]."""
    verts, tris, n_terrain = terrain_arrays(n, light=False)
    rng = np.random.default_rng(seed)
    gx, gz = 100, max(1, n_quads // 100)
    qx, qz = np.meshgrid(np.linspace(-45, 45, gx), np.linspace(-45, 45, gz), indexing="ij")
    qx, qz = qx.reshape(-1)[:n_quads], qz.reshape(-1)[:n_quads]
    qy = 25.0 + rng.uniform(0, 3, n_quads)
    rad = rng.uniform(5, 50, n_quads)
    h = 0.25
    lv = np.stack([np.stack([qx - h, qy, qz - h], -1), np.stack([qx + h, qy, qz - h], -1),
                   np.stack([qx - h, qy, qz + h], -1), np.stack([qx + h, qy, qz + h], -1)], 1).reshape(-1, 3)
    base = verts.shape[0] + 4 * np.arange(n_quads)
    lt = np.stack([np.stack([base, base + 1, base + 2], -1), np.stack([base + 1, base + 3, base + 2], -1)],
                  1).reshape(-1, 3)
    verts = np.concatenate([verts, lv.astype(F32)])
    tris = np.concatenate([tris, lt.astype(np.int32)])
    mats = [DiffuseMaterial((0.7, 0.6, 0.5))] + [EmissiveMaterial((float(r),) * 3) for r in rad]
    mi = np.concatenate([np.zeros(n_terrain, np.int32), np.repeat(1 + np.arange(n_quads), 2).astype(np.int32)])
    mesh = AkariMesh(materials=mats, vertices=verts, indices=tris, material_indices=mi)
    cam = PerspectiveCamera(fov=60, position=(0, 20, 70), rotation=(0, -15, 0), resolution=resolution)
    return Scene(cam, Path(spp=spp, max_depth=max_depth, ray_clamp=0.0, rr_start_depth=rr_start_depth, **path_kw),
                 [mesh])


def mixed_box(resolution=(96, 96), spp=8, max_depth=5, **path_kw):
    """Small closed test scene exercising every reference material kind: Diffuse, Glossy, Mix (nested) and
    Emissive, on Cornell geometry."""
    sc = cornell_box(resolution, spp, max_depth, **path_kw)
    mesh = sc.shapes[0]
    glossy = GlossyMaterial((0.9, 0.9, 0.9), 0.4)
    mix = MixMaterial(0.35, DiffuseMaterial((0.2, 0.3, 0.8)), glossy)
    nested = MixMaterial(0.5, mix, DiffuseMaterial((0.8, 0.8, 0.1)))
    mesh.materials[2] = glossy  # floor
    mesh.materials[5] = mix  # short box
    mesh.materials[6] = nested  # tall box
    return sc


def _checker(w, h, a, b, seed):
    """Procedural test image: two-colour checker with per-texel noise, quantised to bytes like a decoded PNG."""
    rng = np.random.default_rng(seed)
    yy, xx = np.mgrid[0:h, 0:w]
    base = np.where(((xx // 2 + yy // 2) % 2)[..., None] == 0, np.array(a, F32), np.array(b, F32))
    rgb8 = np.clip(np.rint((base + rng.uniform(-0.08, 0.08, base.shape)) * 255.0), 0, 255).astype(np.uint8)
    rgba = np.ones((h, w, 4), F32)
    rgba[..., :3] = rgb8.astype(F32) / F32(255.0)
    return rgba


def textured_box(resolution=(48, 48), spp=8, max_depth=5, **path_kw):
    """Cornell geometry with an image texture in every texture slot the reference has (kernel/texture.h:39-57):
    Diffuse.color, Glossy.color + roughness, Mix.fraction and Emissive.color; texcoords are stretched past [0,1] and
    below 0 so that the fmod / flip / clamp rules of the lookup all matter."""
    from .nodes import ImageTexture
    sc = cornell_box(resolution, spp, max_depth, **path_kw)
    mesh = sc.shapes[0]
    mesh.texcoords = (mesh.texcoords.reshape(-1, 2) * np.array([2.5, 1.75], F32) - np.array([0.75, 0.4], F32)).reshape(-1).astype(F32)
    floor = ImageTexture(rgba=_checker(16, 12, (0.8, 0.75, 0.7), (0.25, 0.3, 0.6), 1))
    wall = ImageTexture(rgba=_checker(8, 8, (0.9, 0.9, 0.9), (0.6, 0.2, 0.2), 2))
    rough = ImageTexture(rgba=_checker(5, 7, (0.2, 0.2, 0.2), (0.7, 0.7, 0.7), 3))
    frac = ImageTexture(rgba=_checker(6, 4, (0.15, 0.5, 0.5), (0.85, 0.5, 0.5), 4))
    glow = ImageTexture(rgba=_checker(4, 4, (1.0, 0.9, 0.6), (0.5, 0.2, 0.1), 5) * np.array([17, 17, 17, 1], F32))
    mesh.materials[2] = DiffuseMaterial(floor)
    mesh.materials[4] = GlossyMaterial(wall, rough)
    mesh.materials[5] = MixMaterial(frac, DiffuseMaterial((0.2, 0.7, 0.3)), GlossyMaterial((0.9, 0.9, 0.9), 0.35))
    mesh.materials[6] = MixMaterial(0.4, DiffuseMaterial(floor), GlossyMaterial(wall, 0.5))
    mesh.materials[7] = EmissiveMaterial(glow)
    return sc
