"""Seeded procedural inputs for the benchmark configs: sorghum plant, field, ground,
incoherent ray set, LiDAR grid, camera.  The reference ships no meshes or scenes, so these are synthetic;
they are shared by the oracle and the GPU path (same arrays go to both).

A scene description mirrors the three maps RayTracerLayer::UpdateMeshesStorage fills
(/root/reference/src/RayTracerFacility/RayTracerLayer.cpp:18-346):
    geometries[handle] = (vertices: VERTEX_DTYPE[n], triangles: uint32[m,3])
    materials[handle]  = structs.Material
    instances[handle]  = (geometry_handle, material_handle, transform float32[16] column-major, component_handle)
"""
import math
from dataclasses import dataclass, field

import numpy as np

from .structs import VERTEX_DTYPE, Material, ScanGrid


@dataclass
class SceneDesc:
    geometries: dict = field(default_factory=dict)
    materials: dict = field(default_factory=dict)
    instances: dict = field(default_factory=dict)
    textures: dict = field(default_factory=dict)           # handle -> (h, w, 4) float32
    material_textures: dict = field(default_factory=dict)  # material -> (albedo, normal, metallic, roughness) handles, 0 = none
    btfs: dict = field(default_factory=dict)               # handle -> structs.BTFTables
    material_btf: dict = field(default_factory=dict)       # material -> btf handle
    env_map: object = None                                 # six texture handles (+X -X +Y -Y +Z -Z) or None

    def triangle_count(self):
        return sum(len(self.geometries[g][1]) for (g, _, _, _) in self.instances.values())


# ---- deterministic counter hash (independent of numpy's RNG implementation) -----------------
def _splitmix64(x):
    x = (x + np.uint64(0x9E3779B97F4A7C15)).astype(np.uint64)
    z = x
    z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
    z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
    return z ^ (z >> np.uint64(31))


def hash_uniform(seed, idx, dim):
    """float64 uniforms in [0,1) from (seed, idx, dim)."""
    with np.errstate(over="ignore"):
        idx = np.asarray(idx, dtype=np.uint64)
        k = _splitmix64(idx * np.uint64(0x100000001B3) + np.uint64(dim) * np.uint64(0x9E3779B1) + np.uint64(seed) *
                        np.uint64(0xD6E8FEB86659FD93))
        k = _splitmix64(k)
    return (k >> np.uint64(11)).astype(np.float64) * (1.0 / 9007199254740992.0)


def _smooth_normals(pos, tris):
    fn = np.cross(pos[tris[:, 1]] - pos[tris[:, 0]], pos[tris[:, 2]] - pos[tris[:, 0]])
    n = np.zeros_like(pos)
    for k in range(3):
        np.add.at(n, tris[:, k], fn)
    ln = np.linalg.norm(n, axis=1, keepdims=True)
    ln[ln == 0] = 1.0
    return n / ln


def make_vertices(pos, nrm, color=(1.0, 1.0, 1.0, 1.0), uv=None, tangent=None):
    v = np.zeros(len(pos), dtype=VERTEX_DTYPE)
    v["position"] = pos.astype(np.float32)
    v["normal"] = nrm.astype(np.float32)
    if tangent is None:
        t = np.cross(nrm, np.array([0.0, 1.0, 0.0]))
        bad = np.linalg.norm(t, axis=1) < 1e-6
        t[bad] = np.cross(nrm[bad], np.array([1.0, 0.0, 0.0]))
        t /= np.linalg.norm(t, axis=1, keepdims=True)
        tangent = t
    v["tangent"] = tangent.astype(np.float32)
    v["color"] = np.asarray(color, dtype=np.float32)
    if uv is not None:
        v["tex_coord"] = uv.astype(np.float32)
    return v


# ---- sorghum plant: exactly 20 000 triangles --------------------------------------------------
def sorghum_plant(seed=1):
    """Stem (12-gon tapered cylinder, 9 rings) + 10 arched ribbon leaves (2 strips x 400 segments) + panicle
    (lat-long ellipsoid 56 x 35).  192 + 16 000 + 3 808 = 20 000 triangles."""
    P, T, UV = [], [], []
    base = 0

    def add(pos, tris, uv):
        nonlocal base
        P.append(pos), T.append(tris + base), UV.append(uv)
        base += len(pos)

    # stem
    height, r0, r1, sides, rings = 1.6, 0.018, 0.008, 12, 9
    ang = np.arange(sides) * (2 * math.pi / sides)
    pos, uv = [], []
    for k in range(rings):
        s = k / (rings - 1)
        r = r0 + (r1 - r0) * s
        pos.append(np.stack([r * np.cos(ang), np.full(sides, height * s), r * np.sin(ang)], 1))
        uv.append(np.stack([np.arange(sides) / sides, np.full(sides, s)], 1))
    pos, uv = np.concatenate(pos), np.concatenate(uv)
    tris = []
    for k in range(rings - 1):
        for j in range(sides):
            a, b = k * sides + j, k * sides + (j + 1) % sides
            c, d = a + sides, b + sides
            tris += [(a, c, b), (b, c, d)]
    add(pos, np.array(tris, dtype=np.int64), uv)
    # leaves
    nseg = 400
    s = np.linspace(0.0, 1.0, nseg + 1)
    for i in range(10):
        u = hash_uniform(seed, i, 0), hash_uniform(seed, i, 1), hash_uniform(seed, i, 2)
        h = 0.25 + 0.12 * i + 0.03 * float(u[0])
        phi = i * math.radians(137.5) + 0.3 * float(u[1])
        length = 0.55 + 0.25 * float(u[2])
        out = np.array([math.cos(phi), 0.0, math.sin(phi)])
        side = np.array([-math.sin(phi), 0.0, math.cos(phi)])
        up = np.array([0.0, 1.0, 0.0])
        p0 = out * 0.012 + up * h
        p1 = p0 + out * (0.45 * length) + up * (0.55 * length)
        p2 = p0 + out * length + up * (0.05 * length)
        rib = ((1 - s) ** 2)[:, None] * p0 + (2 * s * (1 - s))[:, None] * p1 + (s ** 2)[:, None] * p2
        w = 0.045 * np.sin(math.pi * np.clip(s, 1e-4, 1 - 1e-4)) ** 0.7
        fold = 0.35 * w  # slight V-fold
        left = rib - side * w[:, None] + up * fold[:, None]
        right = rib + side * w[:, None] + up * fold[:, None]
        pos = np.empty((3 * (nseg + 1), 3))
        pos[0::3], pos[1::3], pos[2::3] = left, rib, right
        uvl = np.empty((3 * (nseg + 1), 2))
        uvl[0::3] = np.stack([np.zeros_like(s), s], 1)
        uvl[1::3] = np.stack([np.full_like(s, 0.5), s], 1)
        uvl[2::3] = np.stack([np.ones_like(s), s], 1)
        k = np.arange(nseg) * 3
        tris = np.concatenate([
            np.stack([k, k + 3, k + 1], 1), np.stack([k + 1, k + 3, k + 4], 1),
            np.stack([k + 1, k + 4, k + 2], 1), np.stack([k + 2, k + 4, k + 5], 1)])
        add(pos, tris, uvl)
    # panicle
    nlon, nlat = 56, 35
    c = np.array([0.0, height + 0.16, 0.0])
    rad = np.array([0.045, 0.17, 0.045])
    pos, uv = [], []
    for a in range(nlat + 1):
        th = math.pi * a / nlat
        for b in range(nlon):
            ph = 2 * math.pi * b / nlon
            pos.append(c + rad * np.array([math.sin(th) * math.cos(ph), math.cos(th), math.sin(th) * math.sin(ph)]))
            uv.append((b / nlon, a / nlat))
    tris = []
    for a in range(nlat):
        for b in range(nlon):
            i0, i1 = a * nlon + b, a * nlon + (b + 1) % nlon
            i2, i3 = i0 + nlon, i1 + nlon
            if a != 0:
                tris.append((i0, i1, i2))
            if a != nlat - 1:
                tris.append((i1, i3, i2))
    add(np.array(pos), np.array(tris, dtype=np.int64), np.array(uv))
    pos, tris, uv = np.concatenate(P), np.concatenate(T), np.concatenate(UV)
    assert len(tris) == 20000, len(tris)
    nrm = _smooth_normals(pos, tris)
    verts = make_vertices(pos, nrm, color=(0.3, 0.6, 0.2, 1.0), uv=uv)
    return verts, tris.astype(np.uint32)


def ground_plane(half_x=2.0, half_z=2.0, y=0.0):
    pos = np.array([[-half_x, y, -half_z], [half_x, y, -half_z], [half_x, y, half_z], [-half_x, y, half_z]])
    nrm = np.tile(np.array([0.0, 1.0, 0.0]), (4, 1))
    uv = np.array([[0, 0], [1, 0], [1, 1], [0, 1]], dtype=np.float64)
    tris = np.array([[0, 2, 1], [0, 3, 2]], dtype=np.uint32)  # counter-clockwise seen from +y
    return make_vertices(pos, nrm, color=(0.5, 0.4, 0.3, 1.0), uv=uv), tris


def _trs(yaw, scale, tx, ty, tz):
    c, s = math.cos(yaw) * scale, math.sin(yaw) * scale
    m = np.array([[c, 0, s, tx], [0, scale, 0, ty], [-s, 0, c, tz], [0, 0, 0, 1]], dtype=np.float32)
    return np.ascontiguousarray(m.T).reshape(16)  # column-major glm::mat4


IDENTITY = np.eye(4, dtype=np.float32).reshape(16)
PLANT_MATERIAL = dict(albedo=(0.3, 0.6, 0.2), metallic=0.0, roughness=1.0)
GROUND_MATERIAL = dict(albedo=(0.5, 0.4, 0.3), metallic=0.0, roughness=1.0)


def sorghum_scene(n_plants=1, seed=2, ground=True):
    """C1: n_plants=1 (20 002 triangles).  Field: plants on a grid, spacing 0.75 m x 0.25 m, per-plant yaw
    U[0,2pi) and scale U[0.9,1.1].  50 plants -> 1 M triangles (C2), 250 -> 5 M (C4), 500 -> 10 M (C3)."""
    sd = SceneDesc()
    verts, tris = sorghum_plant()
    sd.geometries[1] = (verts, tris)
    sd.materials[1] = Material(**PLANT_MATERIAL)
    if n_plants == 1:
        sd.instances[10] = (1, 1, IDENTITY.copy(), 1010)
        hx = hz = 2.0
    else:
        rows = max(1, int(round(math.sqrt(n_plants / 3.0))))
        per_row = (n_plants + rows - 1) // rows
        idx = np.arange(n_plants)
        yaw = hash_uniform(seed, idx, 0) * 2 * math.pi
        scl = 0.9 + 0.2 * hash_uniform(seed, idx, 1)
        for i in range(n_plants):
            r, c = divmod(i, per_row)
            x = (r - (rows - 1) / 2) * 0.75
            z = (c - (per_row - 1) / 2) * 0.25
            sd.instances[10 + i] = (1, 1, _trs(float(yaw[i]), float(scl[i]), x, 0.0, z), 1010 + i)
        hx = rows * 0.75 / 2 + 1.0
        hz = per_row * 0.25 / 2 + 1.0
    if ground:
        gv, gt = ground_plane(hx, hz)
        sd.geometries[2] = (gv, gt)
        sd.materials[2] = Material(**GROUND_MATERIAL)
        sd.instances[5] = (2, 2, IDENTITY.copy(), 1005)  # lowest handle: ground = primitives 0,1
    return sd


def random_soup(n_tris, seed=7, extent=1.0, size=0.08):
    """Unstructured triangle soup (edge cases for the builder: duplicates, degenerate and tiny triangles)."""
    idx = np.arange(n_tris)
    c = np.stack([hash_uniform(seed, idx, d) for d in range(3)], 1) * extent
    e1 = (np.stack([hash_uniform(seed, idx, 3 + d) for d in range(3)], 1) - 0.5) * size
    e2 = (np.stack([hash_uniform(seed, idx, 6 + d) for d in range(3)], 1) - 0.5) * size
    pos = np.empty((3 * n_tris, 3))
    pos[0::3], pos[1::3], pos[2::3] = c, c + e1, c + e2
    tris = np.arange(3 * n_tris, dtype=np.uint32).reshape(-1, 3)
    nrm = np.repeat(np.cross(e1, e2), 3, axis=0)
    ln = np.linalg.norm(nrm, axis=1, keepdims=True)
    ln[ln == 0] = 1
    sd = SceneDesc()
    sd.geometries[1] = (make_vertices(pos, nrm / ln), tris)
    sd.materials[1] = Material(albedo=(0.7, 0.7, 0.7))
    sd.instances[1] = (1, 1, IDENTITY.copy(), 101)
    return sd


def incoherent_rays(n, lo, hi, seed=3, tmin=0.0, tmax=1e20, start=0):
    """M1 ray set: origins uniform in the box [lo,hi], directions uniform on the sphere.  n x 8 float32."""
    idx = np.arange(start, start + n, dtype=np.uint64)
    lo, hi = np.asarray(lo, dtype=np.float64), np.asarray(hi, dtype=np.float64)
    o = np.stack([lo[d] + (hi[d] - lo[d]) * hash_uniform(seed, idx, d) for d in range(3)], 1)
    z = 1.0 - 2.0 * hash_uniform(seed, idx, 3)
    ph = 2 * math.pi * hash_uniform(seed, idx, 4)
    r = np.sqrt(np.maximum(0.0, 1.0 - z * z))
    rays = np.empty((n, 8), dtype=np.float32)
    rays[:, 0:3] = o
    rays[:, 3] = tmin
    rays[:, 4], rays[:, 5], rays[:, 6] = r * np.cos(ph), z, r * np.sin(ph)
    rays[:, 7] = tmax
    return rays


def lidar_grid(center, size, distance, rotate_angle_deg=0.0, rotation=None):
    """PointCloudScanner::Scan parameters (PointCloudScanner.cpp:70-85).  rotation: 3x3 matrix of the scanner
    (columns = its x,y,z axes); default looks down -Y with up = -Z."""
    if rotation is None:
        rotation = np.array([[1, 0, 0], [0, 0, 1], [0, -1, 0]], dtype=np.float64)  # front (0,0,-1) -> (0,-1,0)
    R = np.asarray(rotation, dtype=np.float64)
    front, up, left = R @ np.array([0.0, 0, -1]), R @ np.array([0.0, 1, 0]), R @ np.array([1.0, 0, 0])
    a = math.radians(rotate_angle_deg)
    # glm::rotate(v, angle, axis): Rodrigues
    k = up / np.linalg.norm(up)
    actual = front * math.cos(a) + np.cross(k, front) * math.sin(a) + k * np.dot(k, front) * (1 - math.cos(a))
    g = ScanGrid()
    g.center[:] = [float(x) for x in center]
    g.left[:] = [float(x) for x in left]
    g.up[:] = [float(x) for x in up]
    g.direction[:] = [float(x) for x in actual]
    g.size[:] = [float(size[0]), float(size[1])]
    g.distance[:] = [float(distance[0]), float(distance[1])]
    return g


def scene_bounds(sd):
    lo, hi = np.full(3, np.inf), np.full(3, -np.inf)
    for (g, _, x, _) in sd.instances.values():
        v = sd.geometries[g][0]["position"].astype(np.float64)
        m = np.asarray(x, dtype=np.float64).reshape(4, 4).T
        w = v @ m[:3, :3].T + m[:3, 3]
        lo, hi = np.minimum(lo, w.min(0)), np.maximum(hi, w.max(0))
    return lo, hi


# ---- device-side generation of the same ray set (bench: 2^26 rays would take ~a minute in numpy) ------------
def _s64(x):
    x &= (1 << 64) - 1
    return x - (1 << 64) if x >= (1 << 63) else x


def _lsr(z, k):
    return (z >> k) & ((1 << (64 - k)) - 1)


def _splitmix64_torch(x):
    x = x + _s64(0x9E3779B97F4A7C15)
    z = x
    z = (z ^ _lsr(z, 30)) * _s64(0xBF58476D1CE4E5B9)
    z = (z ^ _lsr(z, 27)) * _s64(0x94D049BB133111EB)
    return z ^ _lsr(z, 31)


def hash_uniform_torch(seed, idx, dim):
    """hash_uniform() with torch int64 tensors (two's-complement wrap == uint64 arithmetic); float64 out."""
    import torch
    k = idx * _s64(0x100000001B3) + _s64(dim * 0x9E3779B1) + _s64(seed * 0xD6E8FEB86659FD93)
    k = _splitmix64_torch(_splitmix64_torch(k))
    return _lsr(k, 11).to(torch.float64) * (1.0 / 9007199254740992.0)


def incoherent_rays_torch(n, lo, hi, seed=3, tmin=0.0, tmax=1e20, start=0, device="cuda", chunk=1 << 22):
    """incoherent_rays() generated on the device; (n, 8) float32 tensor.  Integer hashing is bit-identical to the
    numpy version; cos/sin(float64) may differ in the last float32 bit on a vanishing fraction of rays."""
    import torch
    out = torch.empty((n, 8), dtype=torch.float32, device=device)
    lo = [float(x) for x in lo]
    hi = [float(x) for x in hi]
    for b in range(0, n, chunk):
        e = min(n, b + chunk)
        idx = torch.arange(start + b, start + e, dtype=torch.int64, device=device)
        o = out[b:e]
        for d in range(3):
            o[:, d] = (lo[d] + (hi[d] - lo[d]) * hash_uniform_torch(seed, idx, d)).to(torch.float32)
        z = 1.0 - 2.0 * hash_uniform_torch(seed, idx, 3)
        ph = (2 * math.pi) * hash_uniform_torch(seed, idx, 4)
        r = torch.sqrt(torch.clamp(1.0 - z * z, min=0.0))
        o[:, 3] = tmin
        o[:, 4] = (r * torch.cos(ph)).to(torch.float32)
        o[:, 5] = z.to(torch.float32)
        o[:, 6] = (r * torch.sin(ph)).to(torch.float32)
        o[:, 7] = tmax
    return out


# ---- synthetic material inputs (the reference ships no textures and no BTF database) -----------------
def checker_texture(width, height, seed=5, cells=8, alpha=None):
    """Smooth-ish seeded RGBA image: a checker modulated by hashed per-cell colours.  (h, w, 4) float32."""
    y, x = np.mgrid[0:height, 0:width]
    cx, cy = (x * cells // width).astype(np.uint64), (y * cells // height).astype(np.uint64)
    cell = cy * np.uint64(cells) + cx
    img = np.stack([hash_uniform(seed, cell, k) for k in range(3)], axis=-1)
    img = 0.15 + 0.8 * img * (0.6 + 0.4 * ((cx + cy) % np.uint64(2))[..., None].astype(np.float64))
    a = np.ones((height, width, 1)) if alpha is None else np.clip(alpha, 0, 1).reshape(height, width, 1)
    return np.concatenate([img, a], axis=-1).astype(np.float32)


def synthetic_btf(seed=4, num_beta=13, num_alpha=7, num_theta=7, num_phi=16, texels=(256, 256), n_pdf1d=4096,
                  n_colors=512, n_index=1024, n_pdf2d_color=512, n_pdf2d_lum=1024, n_pdf2d=2048, n_pdf3d=1024,
                  n_pdf4d=512, **params):
    """Random but valid MLVQ database honouring the loader's table shapes (CompressedBTF.cpp:339-377): every index
    table points inside the table below it; luminance slices are smooth positive curves; scales are near 1."""
    from . import structs as S

    def u(shape, salt):
        n = int(np.prod(shape))
        return hash_uniform(seed, np.arange(n, dtype=np.uint64), salt).reshape(shape)

    def idx(shape, hi, salt, base=0):
        return (np.minimum((u(shape, salt) * hi).astype(np.int64), hi - 1) + base).astype(np.int32)

    beta = np.linspace(-1.0, 1.0, num_beta)
    amp, width = 40.0 + 160.0 * u((n_pdf1d, 1), 1), 0.3 + 1.2 * u((n_pdf1d, 1), 2)
    centre = 0.6 * (u((n_pdf1d, 1), 3) - 0.5)
    pdf1d = amp * np.exp(-((beta[None, :] - centre) / width) ** 2) + 8.0 * u((n_pdf1d, num_beta), 4)
    tables = dict(
        pdf1d_basis=pdf1d,
        colors=np.concatenate([90.0 + 70.0 * u((n_colors, 1), 5), 100.0 + 60.0 * u((n_colors, 1), 6)], axis=1),  # Cb, Cr
        index_ab=idx((n_index, num_beta), n_colors, 7),
        pdf2d_colors=idx((n_pdf2d_color, num_alpha), n_index, 8),
        pdf2d_slices=idx((n_pdf2d_lum, num_alpha), n_pdf1d, 9),
        pdf2d_scales=0.7 + 0.6 * u((n_pdf2d_lum, num_alpha), 10),
        pdf2d_indices=np.stack([idx((n_pdf2d,), n_pdf2d_lum, 11), idx((n_pdf2d,), n_pdf2d_color, 12)], axis=1),
        pdf3d_slices=idx((n_pdf3d, num_theta), n_pdf2d, 13),
        pdf3d_scales=0.8 + 0.4 * u((n_pdf3d, num_theta), 14),
        pdf4d_slices=idx((n_pdf4d, num_phi), n_pdf3d, 15),
        pdf4d_scales=0.8 + 0.4 * u((n_pdf4d, num_phi), 16),
        pdf6_slices=idx(texels, n_pdf4d, 17, base=1),  # 1-based (PDF6D.cuh:88)
        pdf6_scales=0.8 + 0.4 * u(texels, 18),
    )
    return S.BTFTables(tables, **params)


def write_btf_folder(folder, btf, material="synth"):
    """Writes a BTFTables database in the reference's on-disk folder format (all_materialInfo.txt,
    <material>_materialInfo.txt, 13 'rows cols min max' ASCII tables; CompressedBTF.cpp:7-50, 109-228, 284-377)."""
    import os
    os.makedirs(folder, exist_ok=True)
    t, p = btf.t, btf.params
    nb, na = t["pdf1d_basis"].shape[1], t["pdf2d_slices"].shape[1]
    nt, nph = t["pdf3d_slices"].shape[1], t["pdf4d_slices"].shape[1]
    hdr_value = p["hdr_value"] if p["hdr"] else 1.0
    info = f"{material} in out tmp {p['rows_offset']} {p['cols_offset']} {t['pdf6_slices'].shape[0]} {t['pdf6_slices'].shape[1]} {hdr_value:.9g}\n"
    with open(os.path.join(folder, "all_materialInfo.txt"), "w") as f:
        f.write("synthetic_btf_base\n1\n1\n1\n0\n0\n")          # name, loadMaterials, maxMaterials, allInOne, 34view, compact
        f.write("3\n81\n81\n256\n")                              # ncolour nview nillu tileSize
        f.write(f"{p['use_cos_beta']} {nb} 0 {na} 0 {nt} 0 {nph} 0 0 0 {p['code_btf_flag']}\n")
        f.write("0 0 0 0 0 0 0 0 0 0\n0 0 0 0 0 0 0 0 0 0\n")
        f.write("0 0.1 0.1 0.1 0.1 0.1 0.1 0.1 0.1 0.1 0.1 0.1\n")
        f.write(info)
    with open(os.path.join(folder, f"{material}_materialInfo.txt"), "w") as f:
        f.write(info)
    names = {"pdf6_slices": f"{material}_PDF6Dslices.txt", "pdf6_scales": f"{material}_PDF6Dscale.txt",
             "pdf1d_basis": "all_PDF1Dslice.txt", "colors": "all_colors.txt", "index_ab": "all_indexAB.txt",
             "pdf2d_colors": "all_PDF2Dcolours.txt", "pdf2d_slices": "all_PDF2Dslices.txt",
             "pdf2d_scales": "all_PDF2Dscale.txt", "pdf2d_indices": "all_PDF2Dindices.txt",
             "pdf3d_scales": "all_PDF3Dscale.txt", "pdf3d_slices": "all_PDF3Dslices.txt",
             "pdf4d_scales": "all_PDF4Dscale.txt", "pdf4d_slices": "all_PDF4Dslices.txt"}
    for k, fn in names.items():
        a = t[k]
        with open(os.path.join(folder, fn), "w") as f:
            if a.dtype.kind == "i":
                f.write(f"{a.shape[0]} {a.shape[1]} {int(a.min())} {int(a.max())}\n")
                for row in a:
                    f.write(" ".join(str(int(v)) for v in row) + " \n")
            else:
                f.write(f"{a.shape[0]} {a.shape[1]} {float(a.min()):.9g} {float(a.max()):.9g}\n")
                for row in a:
                    f.write(" ".join(f"{float(v):.9g}" for v in row) + " \n")


def _quad(cx, cy, cz, ex, ez, tilt_deg=0.0, uv_scale=1.0, n=4, seed=0):
    """n x n grid quad centred at (cx,cy,cz), half extents (ex, ez), tilted about x; uv in [0, uv_scale]; per-vertex
    hashed colours; analytic tangent = d(position)/du."""
    g = np.linspace(-1.0, 1.0, n + 1)
    u, v = np.meshgrid(g, g, indexing="xy")
    t = math.radians(tilt_deg)
    px, py, pz = u * ex, -v * ez * math.sin(t), v * ez * math.cos(t)
    pos = np.stack([px.ravel() + cx, py.ravel() + cy, pz.ravel() + cz], axis=1)
    nrm = np.tile(np.array([0.0, math.cos(t), math.sin(t)]), (len(pos), 1))
    tan = np.tile(np.array([1.0, 0.0, 0.0]), (len(pos), 1))
    uv = np.stack([(u.ravel() + 1) / 2 * uv_scale, (v.ravel() + 1) / 2 * uv_scale], axis=1)
    idx = np.arange(len(pos), dtype=np.uint64)
    col = np.stack([hash_uniform(seed, idx, k) for k in range(3)] + [np.ones(len(pos))], axis=1)
    verts = make_vertices(pos, nrm, uv=uv, tangent=tan)
    verts["color"] = col.astype(np.float32)
    tris = []
    for j in range(n):
        for i in range(n):
            a = j * (n + 1) + i
            tris += [[a, a + n + 2, a + 1], [a, a + n + 1, a + n + 2]]
    return verts, np.array(tris, dtype=np.uint32)


def materials_scene(seed=9, btf=None):
    """Small scene exercising every material input: textured Default (albedo + normal + metallic + roughness maps, uv
    beyond [0,1] for the wrap mode), VertexColor, CompressedBTF, and a Default material whose albedo texture carries
    alpha < 1 (stochastic any-hit).  Textures 1..5, BTF 1, cube map faces 11..16."""
    from . import structs as S
    sd = SceneDesc()
    sd.geometries[1] = _quad(0.0, 0.0, 0.0, 1.5, 1.5, uv_scale=3.0, n=6, seed=seed)           # ground
    sd.geometries[2] = _quad(-0.5, 0.45, 0.1, 0.45, 0.45, tilt_deg=25.0, n=5, seed=seed + 1)    # vertex colours
    sd.geometries[3] = _quad(0.55, 0.5, -0.2, 0.5, 0.5, tilt_deg=-20.0, uv_scale=2.0, n=5, seed=seed + 2)  # BTF
    sd.geometries[4] = _quad(0.0, 0.95, 0.0, 0.9, 0.9, uv_scale=1.0, n=3, seed=seed + 3)       # alpha-textured canopy
    sd.materials[1] = Material(albedo=(0.8, 0.8, 0.8), metallic=0.2, roughness=0.7)
    sd.materials[2] = Material(albedo=(0.1, 0.1, 0.1), metallic=0.0, roughness=1.0, material_type=S.MAT_VERTEX_COLOR)
    sd.materials[3] = Material(material_type=S.MAT_COMPRESSED_BTF)
    sd.materials[4] = Material(albedo=(0.2, 0.7, 0.3), metallic=0.0, roughness=0.9, emission=0.05)
    for k in (1, 2, 3, 4):
        sd.instances[k] = (k, k, IDENTITY.copy(), 100 + k)
    sd.textures[1] = checker_texture(32, 16, seed=seed)
    nm = checker_texture(16, 16, seed=seed + 1)
    nm[..., 0:2] = 0.5 + 0.3 * (nm[..., 0:2] - 0.5)
    nm[..., 2] = 0.9
    sd.textures[2] = nm
    sd.textures[3] = checker_texture(8, 8, seed=seed + 2, cells=4)
    sd.textures[4] = checker_texture(8, 4, seed=seed + 3, cells=4)
    yy, xx = np.mgrid[0:16, 0:16]
    alpha = 0.15 + 0.8 * hash_uniform(seed + 4, (yy * 16 + xx).astype(np.uint64), 0)
    sd.textures[5] = checker_texture(16, 16, seed=seed + 5, alpha=alpha)
    sd.material_textures[1] = (1, 2, 3, 4)
    sd.material_textures[4] = (5, 0, 0, 0)
    sd.btfs[1] = btf if btf is not None else synthetic_btf(seed=4, texels=(32, 32), n_pdf1d=256, n_colors=64, n_index=128,
                                                            n_pdf2d_color=64, n_pdf2d_lum=128, n_pdf2d=256, n_pdf3d=128,
                                                            n_pdf4d=64)
    sd.material_btf[3] = 1
    for k in range(6):
        sd.textures[11 + k] = checker_texture(8, 8, seed=seed + 20 + k, cells=4)
    sd.env_map = [11, 12, 13, 14, 15, 16]
    return sd
