"""The C ABI: the library loads, exports every symbol include/rtf/rtf.h declares, and the reference-layout
records have the sizes/offsets the reference's structs have (no compute calls: runs without a GPU)."""
import ctypes as C
import os
import re
import subprocess
import tempfile

from .conftest import ROOT


def declared_symbols():
    hdr = open(os.path.join(ROOT, "include", "rtf", "rtf.h")).read()
    return sorted(set(re.findall(r"RTF_API\s+[\w\s\*]+?\b(rtf_\w+)\s*\(", hdr)))


def test_library_exports_every_declared_symbol(mod):
    names = declared_symbols()
    assert len(names) >= 40
    L = mod.lib()
    for n in names:
        assert hasattr(L, n), f"librtf_b200.so does not export {n}"
    assert sorted(mod.EXPORTS) == names
    assert b"sm_100a" in L.rtf_version()


def test_no_gpu_means_loud_failure(mod):
    import torch
    if torch.cuda.is_available():
        return
    h = C.c_void_p()
    rc = mod.lib().rtf_create(0, C.byref(h))
    assert rc == mod.ERR_CUDA and b"no CPU fallback" in mod.lib().rtf_last_error()


def test_struct_layouts_match_header(S):
    src = r'''
#include <stdio.h>
#include <stddef.h>
#include "rtf/rtf.h"
int main(void) {
  printf("%zu %zu %zu %zu %zu %zu %zu %zu %zu\n", sizeof(rtf_vertex), sizeof(rtf_illum_sampler), sizeof(rtf_hit_info),
         sizeof(rtf_pc_sample), sizeof(rtf_env), sizeof(rtf_ray_props), sizeof(rtf_material), sizeof(rtf_camera_props),
         sizeof(rtf_scan_grid));
  printf("%zu %zu %zu %zu %zu %zu\n", offsetof(rtf_illum_sampler, direction), offsetof(rtf_illum_sampler, energy),
         offsetof(rtf_illum_sampler, front_face), offsetof(rtf_pc_sample, handle), offsetof(rtf_pc_sample, hit),
         offsetof(rtf_pc_sample, hit_info));
  return 0;
}'''
    with tempfile.TemporaryDirectory() as d:
        open(os.path.join(d, "t.c"), "w").write(src)
        subprocess.check_call(["gcc", "-std=c99", "-I", os.path.join(ROOT, "include"), "-o", os.path.join(d, "t"), os.path.join(d, "t.c")])
        out = subprocess.check_output([os.path.join(d, "t")]).decode().split()
    sizes = [int(x) for x in out]
    assert sizes[:9] == [C.sizeof(S.Vertex), C.sizeof(S.IlluminationSampler), C.sizeof(S.HitInfo),
                         C.sizeof(S.PointCloudSample), C.sizeof(S.EnvironmentProperties), C.sizeof(S.RayProperties),
                         C.sizeof(S.Material), C.sizeof(S.CameraProps), C.sizeof(S.ScanGrid)]
    # reference sizes: Vertex 80, IlluminationSampler 268 (RayTracer.hpp:187-211), HitInfo 76, PointCloudSample 112
    assert sizes[:4] == [80, 268, 76, 112]
    assert sizes[9:] == [S.IlluminationSampler.direction.offset, S.IlluminationSampler.energy.offset,
                         S.IlluminationSampler.front_face.offset, S.PointCloudSample.handle.offset,
                         S.PointCloudSample.hit.offset, S.PointCloudSample.hit_info.offset]
    assert sizes[9:] == [240, 252, 264, 24, 32, 36]


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "raytracerfacility_b200")
    for dp, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                txt = open(os.path.join(dp, f), errors="ignore").read()
                assert "pyoracle" not in txt and "liboracle" not in txt and "oracle_" not in txt, f


def test_header_is_plain_c_and_cxx(tmp_path):
    """include/rtf/rtf.h is the drop-in boundary: it must compile as C99 and as C++11 with nothing but the standard
    headers (no torch, no CUDA types in the signatures)."""
    import subprocess
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    src = tmp_path / "t.c"
    src.write_text("#include <rtf/rtf.h>\nint main(void) { rtf_env e; (void)e; return (int)sizeof(rtf_illum_sampler) - 268; }\n")
    inc = os.path.join(root, "include")
    subprocess.check_call(["gcc", "-std=c99", "-Wall", "-Wextra", "-pedantic", "-Werror", "-I", inc, "-fsyntax-only", str(src)])
    subprocess.check_call(["g++", "-std=c++11", "-Wall", "-Wextra", "-pedantic", "-Werror", "-I", inc, "-fsyntax-only", "-x", "c++",
                           str(src)])
    text = open(os.path.join(inc, "rtf", "rtf.h")).read()
    includes = re.findall(r"#\s*include\s*[<\"]([^>\"]+)", text)
    assert set(includes) <= {"stddef.h", "stdint.h"}, includes
