"""CPU tests of the boundary: the C-ABI library loads, exports every symbol the
header declares, its POD layouts match the shader's structs, the host-side
constructors reproduce the shader's hard-coded data, and (no GPU here) render
calls fail loudly instead of falling back to a CPU path."""
import ctypes as C
import os
import re

import numpy as np
import pytest

import openglraytracer_b200 as rt
import oracle_binding as ob
from openglraytracer_b200 import _abi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _has_gpu():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


def test_header_symbols_all_exported():
    hdr = open(os.path.join(ROOT, "include", "rt_b200.h")).read()
    declared = set(re.findall(r"^\s*(?:int|void|uint64_t|const char\*|const rt_scene\*|const rt_camera\*)\s+(rt_[a-z0-9_]+)\s*\(", hdr, re.M))
    assert declared, "no declarations parsed"
    assert declared == set(_abi.EXPORTS), declared ^ set(_abi.EXPORTS)
    L = _abi.lib()
    for name in declared:
        assert getattr(L, name) is not None
    assert L.rt_abi_version() == 1


def test_pod_layouts_match_shader_structs():
    # glsl:56-69 Material 20 words, glsl:244-258 Object 33, glsl:190-196 Light 15, glsl:36-50 Camera 10
    assert C.sizeof(rt.Material) == 80 and C.sizeof(rt.Object) == 132
    assert C.sizeof(rt.Light) == 60 and C.sizeof(rt.Camera) == 40
    assert C.sizeof(rt.Material) == C.sizeof(ob.Material) and C.sizeof(rt.Object) == C.sizeof(ob.Object)


def test_ctypes_mirrors_match_the_header(tmp_path):
    """The ctypes mirrors of rt_stats / rt_options / the scene structs have the sizes (and, for the last field,
    the offsets) the C compiler gives the declarations in include/rt_b200.h."""
    import subprocess
    from openglraytracer_b200 import _abi
    src = tmp_path / "sizes.c"
    src.write_text('#include <stdio.h>\n#include <stddef.h>\n#include "rt_b200.h"\n'
                   'int main(void) { printf("%zu %zu %zu %zu %zu %zu %zu %zu %zu %zu\\n", sizeof(rt_stats), sizeof(rt_options), '
                   'sizeof(rt_object), sizeof(rt_light), sizeof(rt_camera), sizeof(rt_material), '
                   'offsetof(rt_stats, shadow_group_passes), offsetof(rt_options, no_cull), offsetof(rt_stats, shaded_lights), '
                   'offsetof(rt_options, done_value)); return 0; }\n')
    exe = tmp_path / "sizes"
    subprocess.check_call(["gcc", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)])
    got = [int(x) for x in subprocess.check_output([str(exe)]).split()]
    want = [C.sizeof(_abi.Stats), C.sizeof(_abi.Options), C.sizeof(rt.Object), C.sizeof(rt.Light), C.sizeof(rt.Camera),
            C.sizeof(rt.Material), _abi.Stats.shadow_group_passes.offset, _abi.Options.no_cull.offset,
            _abi.Stats.shaded_lights.offset, _abi.Options.done_value.offset]
    assert got == want


@pytest.mark.parametrize("time", [0.0, 0.016, 2.5, 7.123, 100.0])
def test_builtin_scene_and_camera_match_oracle_restatement(time):
    a, b = rt.builtin_scene(time), ob.builtin_scene(time)
    assert bytes(a.objects) == bytes(b.objects) and bytes(a.lights) == bytes(b.lights)
    assert a.scene.objects_count == 5 and a.scene.lights_count == 3 and a.scene.max_depth == 0   # glsl:321,224,22
    assert bytes(rt.builtin_camera(time)) == bytes(ob.builtin_camera(time))


def test_builtin_scene_values():
    s = rt.builtin_scene(0.0)
    o = s.objects
    assert list(o[0].box_mins) == [-11.0] * 3 and list(o[0].box_maxs) == [11.0] * 3      # glsl:266-267
    assert o[0].sphere_radius == -1.0 and o[4].sphere_radius == 2.0                     # glsl:179, 314
    assert list(o[4].box_mins) == [0.0] * 3 and list(o[4].box_maxs) == [0.0] * 3        # null_box glsl:178
    assert list(o[1].box_maxs) == [1.5] * 3                                             # 0.5*sin(0)+1.5
    assert list(o[2].angles) == [0.0, 45.0, 0.0] and list(o[3].angles) == [45.0, 0.0, 45.0]
    assert abs(o[4].material.reflectivity - 0.8) < 1e-7 and abs(o[2].material.transparency - 0.6) < 1e-7
    assert abs(o[0].material.diffuse[3] - 0.4) < 1e-7 and o[1].material.refraction_index == 1.0
    assert list(s.lights[2].diffuse) == [1.0, 0.0, 0.0, 1.0] and abs(s.lights[0].ambient[0] - 0.3) < 1e-7


def test_scaled_scene_generator():
    a, b = rt.scaled_scene(1024), rt.scaled_scene(1024)
    assert bytes(a.objects) == bytes(b.objects)                      # deterministic
    assert bytes(rt.scaled_scene(1024, seed=1).objects) != bytes(a.objects)
    o = a.objects
    assert list(o[0].box_maxs) == [11.0] * 3                         # object 0 = enclosing cube
    s = 4.0 * 1024 ** (-1.0 / 3.0)
    n_sph = 0
    for i in range(1, 1024):
        assert all(-9.0 <= p <= 9.0 for p in o[i].position)
        if i % 5 == 4:
            n_sph += 1
            assert list(o[i].box_maxs) == [0.0] * 3 and 0.5 * s <= o[i].sphere_radius <= 1.5 * s
        else:
            assert o[i].sphere_radius == -1.0
            assert all(0.5 * s <= h <= 1.5 * s for h in o[i].box_maxs)
            assert [-m for m in o[i].box_mins] == list(o[i].box_maxs)
    assert n_sph == 204
    # prefix property: the first objects do not depend on n
    assert bytes(rt.scaled_scene(64).objects)[132:] != bytes(a.objects)[132:64 * 132]  # scale s depends on n


def test_turntable_cameras():
    cams = rt.turntable_cameras(256)
    assert len(cams) == 256
    assert bytes(cams[0]) == bytes(rt.orbit_camera(0.5)) == bytes(rt.builtin_camera(0.0))
    r = [np.hypot(c.position[0], c.position[1]) for c in cams]
    assert np.allclose(r, 10.0, atol=1e-5)


def test_ppm_and_raw_writers(tmp_path):
    img = np.zeros((3, 2, 4), np.uint8)
    img[0, :, 0] = 255       # row 0 = bottom scanline -> last in the file
    img[2, :, 2] = 200
    img[..., 3] = 77         # alpha is dropped (frag:9)
    p = str(tmp_path / "a.ppm")
    rt.write_ppm(p, img)
    data = open(p, "rb").read()
    assert data.startswith(b"P6\n2 3\n255\n")
    body = np.frombuffer(data[len(b"P6\n2 3\n255\n"):], np.uint8).reshape(3, 2, 3)
    assert np.array_equal(body, img[::-1, :, :3])
    q = str(tmp_path / "a.raw")
    rt.write_raw(q, img)
    assert open(q, "rb").read() == img.tobytes()


def test_options_default_and_bad_arguments():
    o = rt.Options()
    _abi.lib().rt_options_default(C.byref(o))
    assert o.spp == 1 and o.seed == 0x5EED0001 and o.device == -1 and o.strict == 0 and o.y1 == 0
    L = _abi.lib()
    assert L.rt_create(None, 0) == _abi.RT_ERR_BAD_ARG
    assert b"NULL" in L.rt_last_error_string()
    assert L.rt_builtin_scene(C.c_float(0), None, None, None) == _abi.RT_ERR_BAD_ARG
    assert L.rt_write_ppm(b"/nonexistent-dir/x.ppm", None, 1, 1) == _abi.RT_ERR_BAD_ARG


@pytest.mark.skipif(_has_gpu(), reason="checks the no-GPU failure mode")
def test_no_gpu_fails_loudly_no_cpu_fallback():
    with pytest.raises(rt.RtError) as e:
        rt.render(rt.builtin_scene(), rt.builtin_camera(), 16, 9)
    assert e.value.code == _abi.RT_ERR_NO_DEVICE and "no CPU fallback" in str(e.value)
    with pytest.raises(rt.RtError):
        rt.Raytracer()


def test_product_does_not_reference_oracle():
    pkg = os.path.join(ROOT, "openglraytracer_b200")
    for dirpath, _, files in os.walk(pkg):
        if "build" in dirpath:
            continue
        for f in files:
            if f.endswith((".py", ".cpp", ".cu", ".cuh", ".h", ".inc")):
                src = open(os.path.join(dirpath, f), errors="ignore").read()
                assert "rt_oracle" not in src and "oracle_binding" not in src and "librt_oracle" not in src, f
