"""C-ABI shape: the library loads, exports every symbol include/crt_api.h declares, and refuses to
compute without a GPU (no CPU fallback). No compute calls here."""
import ctypes as C
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_functions():
    text = open(os.path.join(ROOT, "include", "crt_api.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(crt_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol(crt):
    lib = crt.load_library()
    names = declared_functions()
    assert len(names) >= 25
    for name in names:
        assert hasattr(lib, name), name
    assert sorted(crt.API_FUNCTIONS) == names
    assert lib.crt_api_version() == 2


def test_struct_sizes_match_header(crt):
    assert C.sizeof(crt.Material) == 36          # == sizeof(CRT::MaterialData), Material.cuh:16-47
    assert C.sizeof(crt.Camera) == 48
    assert C.sizeof(crt.Config) == 48
    assert C.sizeof(crt.RenderParams) == 32
    assert C.sizeof(crt.Stats) == 120


def test_no_cpu_fallback(crt, gpu_available):
    if gpu_available:
        pytest.skip("GPU present")
    with pytest.raises(crt.CrtError) as e:
        crt.Renderer(64, 64)
    assert e.value.code == crt.E_NO_DEVICE
    assert "no CPU fallback" in str(e.value)


def test_create_rejects_bad_arguments(crt):
    lib = crt.load_library()
    ctx = C.c_void_p()
    cfg = crt.Config(width=0, height=10, device=0, mode=0)
    assert lib.crt_create(C.byref(ctx), C.byref(cfg)) == crt.E_INVALID
    cfg = crt.Config(width=8, height=8, device=0, mode=7)
    assert lib.crt_create(C.byref(ctx), C.byref(cfg)) == crt.E_INVALID
    assert lib.crt_create(None, None) == crt.E_INVALID
    lib.crt_destroy(None)   # no-op


@pytest.mark.parametrize("name,tris,spheres,meshes", [("cornell", 32, 2, 1), ("cornell_blob", 32 + 69632, 2, 2), ("spheres", 0, None, 0)])
def test_builtin_scenes(crt, name, tris, spheres, meshes):
    s = crt.builtin_scene(name)
    assert s.num_triangles == tris
    if spheres is not None:
        assert s.num_spheres == spheres
    else:
        assert 470 <= s.num_spheres <= 500      # reference limit s_MAX_OBJECTS = 500 (HittableList.cuh:70)
        assert s.num_materials == s.num_spheres
    assert s.num_meshes == meshes
    a = s.arrays()
    if tris:
        assert a["indices"].max() < s.num_vertices
        assert a["tri_material"].min() >= 0 and a["tri_material"].max() < s.num_materials
        # reference units: max extent 0.6 about the origin (SceneManager.h:307-325)
        assert np.abs(a["positions"]).max() <= 0.3 + 1e-6
    s.free()


def test_builtin_terrain_is_scalable(crt):
    s = crt.builtin_scene("terrain", 40)
    assert s.num_triangles == 12 + 40 * 39 * 2
    s.free()
    with pytest.raises(crt.CrtError):
        crt.builtin_scene("no_such_scene")


def test_obj_loader_reference_rules(crt, tmp_path):
    # a quad (fan-triangulated) with an emissive and a metal material + a second, MTL-less file
    (tmp_path / "a.mtl").write_text("newmtl lamp\nKe 5 4 3\nKd 0 0 0\nnewmtl steel\nKd 0.7 0.6 0.5\nKs 0.5 0.5 0.5\nNs 98\n"
                                    "newmtl glass\nKd 1 1 1\nd 0.5\nNi 1.5\nnewmtl wall\nKd 0.2 0.3 0.4\n")
    (tmp_path / "a.obj").write_text("mtllib a.mtl\nv 0 0 0\nv 2 0 0\nv 2 2 0\nv 0 2 0\nv 0 0 4\nusemtl lamp\nf 1 2 3 4\n"
                                    "usemtl steel\nf 1/1/1 2/2/2 5/3/3\nusemtl glass\nf -1 -2 -3\nusemtl wall\nf 1//1 3//1 5//1\n")
    (tmp_path / "b.obj").write_text("v 0 0 0\nv 1 0 0\nv 0 1 0\nf 1 2 3\n")
    s = crt.load_obj([str(tmp_path / "a.obj"), str(tmp_path / "b.obj")])
    a = s.arrays()
    assert s.num_triangles == 6 and s.num_meshes == 2 and s.num_spheres == 2
    mats = a["materials"]
    assert [m[0] for m in mats] == [crt.MAT_DIFFUSE_LIGHT, crt.MAT_METAL, crt.MAT_DIELECTRIC, crt.MAT_LAMBERTIAN,
                                    crt.MAT_LAMBERTIAN, crt.MAT_METAL]
    assert mats[1][3] == pytest.approx(np.sqrt(np.float32(2.0) / np.float32(100.0)), rel=1e-6)   # sqrt(2/(Ns+2))
    assert mats[2][4] == pytest.approx(1.5)
    assert list(a["tri_material"][:5]) == [0, 0, 1, 2, 3]
    # mesh 1's ids are offset by the number of DISTINCT ids mesh 0 used (4) -> index 4 = ground Lambertian
    assert a["tri_material"][5] == 4
    # vertex arrays are 3x over-allocated per file with a zero tail (SceneManager.h:253)
    assert s.num_vertices == 3 * 5 + 3 * 3
    # everything re-normalised to a 0.6 max extent
    ext = a["positions"].max(axis=0) - a["positions"].min(axis=0)
    assert ext.max() == pytest.approx(0.6, rel=1e-5)
    s.free()
    with pytest.raises(crt.CrtError) as e:
        crt.load_obj([str(tmp_path / "missing.obj")])
    assert e.value.code == crt.E_IO
    (tmp_path / "bad.obj").write_text("v 0 0 0\nf 1 2 3\n")
    with pytest.raises(crt.CrtError) as e:
        crt.load_obj([str(tmp_path / "bad.obj")])
    assert e.value.code == crt.E_PARSE
