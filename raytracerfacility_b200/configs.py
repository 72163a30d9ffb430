"""The five BASELINE.json configurations as data: scene, environment, ray properties, camera pose / scan grid.
One place, so that bench.py and tests/test_gpu_configs.py run the SAME jobs (SURVEY.md 8(d)
"Synthetic inputs": the reference ships no meshes, scenes or BTF data, everything is generated from fixed seeds).

C1  per-triangle illumination, one sorghum plant + ground (20 002 triangles), sky light, 64 samples/triangle
C2  camera 1920x1080, 16 spp, 4 bounces, 1 000 002-triangle field; brdf=lambert_ggx (roughness 0.5) and brdf=reference
C3  per-triangle illumination of the 10 000 002-triangle field, 256 samples/triangle
C4  virtual LiDAR, 10 000 x 10 000 rays from 30 m above the 5 000 002-triangle field, looking -Y
C5  compressed-BTF camera render 3840x2160, 64 spp (synthetic MLVQ database, seed 4)
"""
from dataclasses import dataclass, field

from . import scenes, structs as S

SEED = 0x5EED
PUSH = 1e-3
CAMERA_POSITION = (0.0, 1.5, 4.0)
CAMERA_ROTATION_WXYZ = (0.9950042, -0.0998334, 0.0, 0.0)  # slight downward pitch towards the origin
PLANTS = {"C1": 1, "C2": 50, "C3": 500, "C4": 250, "C5": 50}


@dataclass
class Config:
    name: str
    what: str
    scene: object
    env: object = field(default_factory=S.EnvironmentProperties)
    rp: object = None
    width: int = 0
    height: int = 0
    grid: object = None
    brdf: int = S.BRDF_REFERENCE

    @property
    def triangles(self):
        return self.scene.triangle_count()


def c1():
    return Config("C1", "per-triangle illumination, 20 002 triangles x 64 samples x 2 faces, 4 bounces, sky light, seed 0x5EED",
                  scenes.sorghum_scene(PLANTS["C1"]), rp=S.RayProperties(bounces=4, samples=64))


def c2(brdf=S.BRDF_LAMBERT_GGX):
    sd = scenes.sorghum_scene(PLANTS["C2"])
    if brdf == S.BRDF_LAMBERT_GGX:  # C2 brdf=lambert_ggx runs with roughness 0.5
        for mh, mat in (1, scenes.PLANT_MATERIAL), (2, scenes.GROUND_MATERIAL):
            sd.materials[mh] = S.Material(**dict(mat, roughness=0.5))
    name = "lambert_ggx, roughness 0.5" if brdf == S.BRDF_LAMBERT_GGX else "reference BRDF"
    return Config("C2", f"camera 1920x1080, 16 spp, 4 bounces, {name}, 1 000 002 triangles", sd,
                  rp=S.RayProperties(bounces=4, samples=16), width=1920, height=1080, brdf=brdf)


def c3(samples=256):
    return Config("C3", f"per-triangle illumination, 10 000 002 triangles x {samples} samples x 2 faces, 4 bounces, sky light",
                  scenes.sorghum_scene(PLANTS["C3"]), rp=S.RayProperties(bounces=4, samples=samples))


def c4(side=10000):
    sd = scenes.sorghum_scene(PLANTS["C4"])
    lo, hi = scenes.scene_bounds(sd)
    ext = hi - lo
    grid = scenes.lidar_grid(center=((lo[0] + hi[0]) / 2, 30.0, (lo[2] + hi[2]) / 2), size=(float(ext[0]), float(ext[2])),
                             distance=(float(ext[0]) / side, float(ext[2]) / side))
    return Config("C4", f"virtual LiDAR, {side} x {side} rays from 30 m above the 5 000 002-triangle field", sd, grid=grid)


def c5(spp=64, width=3840, height=2160):
    sd = scenes.sorghum_scene(PLANTS["C5"])
    sd.materials[1] = S.Material(material_type=S.MAT_COMPRESSED_BTF)
    sd.btfs[1] = scenes.synthetic_btf(seed=4, multiplier=0.0487)
    sd.material_btf[1] = 1
    return Config("C5", f"compressed-BTF camera {width}x{height}, {spp} spp, 4 bounces, synthetic MLVQ database "
                  "(13x7x7x16, 256x256 texels)", sd, rp=S.RayProperties(bounces=4, samples=spp), width=width, height=height)


def setup_camera(cam, cfg):
    """Pose and size of the C2 / C5 camera (fov 120 -> 60 degrees effective, aperture 0, no accumulation)."""
    cam.resize(cfg.width, cfg.height)
    cam.props.accumulate = 0
    cam.set(CAMERA_POSITION, CAMERA_ROTATION_WXYZ)
    return cam


def clear_scene(rt, sd):
    """Remove what `sd` put into the context, so that the next configuration starts from an empty scene."""
    for ih in sd.instances:
        rt.remove_instance(ih)
    for gh in sd.geometries:
        rt.remove_geometry(gh)
    for mh in sd.materials:
        rt.remove_material(mh)
    for bh in sd.btfs:
        rt.remove_btf(bh)
