"""Scene / camera description on the wire: `rtscene 1` text files.
Host logic only: runs without a GPU."""
import ctypes as C

import numpy as np
import pytest

import openglraytracer_b200 as rt


def _bytes(x):
    return bytes(memoryview(x).cast("B"))


def test_builtin_scene_round_trips_bit_exactly(tmp_path):
    for t in (0.0, 2.5, 3.7):
        sc, cam = rt.builtin_scene(t, 4), rt.builtin_camera(t)
        p = str(tmp_path / "builtin.rtscene")
        rt.save_scene(p, sc, cam)
        sc2, cam2 = rt.load_scene(p)
        assert sc2.objects_count == 5 and sc2.scene.lights_count == 3 and sc2.max_depth == 4
        assert _bytes(sc2.objects)[: 5 * 132] == _bytes(sc.objects)
        assert _bytes(sc2.lights)[: 3 * 60] == _bytes(sc.lights)
        assert _bytes(cam2) == _bytes(cam)


def test_scaled_scene_round_trips_bit_exactly(tmp_path):
    sc = rt.scaled_scene(1024, max_depth=2)
    p = str(tmp_path / "scaled.rtscene")
    rt.save_scene(p, sc)                       # no camera record
    sc2, cam2 = rt.load_scene(p)
    assert cam2 is None and sc2.objects_count == 1024 and sc2.max_depth == 2
    assert _bytes(sc2.objects) == _bytes(sc.objects)


def test_hand_written_file_with_comments_and_nulls(tmp_path):
    p = tmp_path / "hand.rtscene"
    mat = "1 1 1 1  0 1 0 1  1 1 1 1  10  0 0 0 0  0.4 0.6 1.5"
    p.write_text("# a comment\nrtscene 1\n\ndepth 3   # trailing comment\n"
                 "camera 10 0 0  0 90 0  90 1.7777778 0.1 1000\n"
                 "light 7 7 2  0.05 0.05 0.05 0.05  1 1 1 1  1 1 1 1\n"
                 "object 0 0 0 0 0 0  2  0 0 0  0 0 0  " + mat + "\n"          # sphere (null box)
                 "object -1 -1 -1 1 1 1  -1  3 0 0  0 45 0  " + mat + "\n")    # box (null sphere)
    sc, cam = rt.load_scene(str(p))
    assert sc.max_depth == 3 and sc.objects_count == 2 and sc.scene.lights_count == 1
    assert sc.objects[0].sphere_radius == 2.0 and list(sc.objects[0].box_maxs) == [0.0, 0.0, 0.0]
    assert sc.objects[1].sphere_radius == -1.0 and sc.objects[1].angles[1] == 45.0
    assert sc.objects[1].material.refraction_index == 1.5 and sc.objects[1].material.shininess == 10.0
    assert cam.v_fov == 90.0 and np.float32(cam.aspect) == np.float32(1.7777778)


@pytest.mark.parametrize("text,why", [
    ("", "empty"), ("rtscene 2\n", "header"), ("depth 1\n", "header"),
    ("rtscene 1\nobject 1 2 3\n", "33 numbers"), ("rtscene 1\nlight 1 2 3 4\n", "15 numbers"),
    ("rtscene 1\ncamera 1 2 3 4 5 6 7 8 9 10 11\n", "trailing"), ("rtscene 1\nteapot 1\n", "unknown"),
    ("rtscene 1\ndepth -1\n", "depth"), ("rtscene 1\ndepth x\n", "depth"),
])
def test_malformed_files_are_rejected_with_a_message(tmp_path, text, why):
    p = tmp_path / "bad.rtscene"
    p.write_text(text)
    with pytest.raises(rt.RtError) as e:
        rt.load_scene(str(p))
    assert e.value.code == -1 and why in str(e.value)


def test_missing_file_and_unwritable_path():
    with pytest.raises(rt.RtError):
        rt.load_scene("/nonexistent/dir/x.rtscene")
    with pytest.raises(rt.RtError):
        rt.save_scene("/nonexistent/dir/x.rtscene", rt.builtin_scene())


def _cli():
    import os
    return os.path.join(os.path.dirname(rt.__file__), "rt_cli")


def test_cli_fails_loudly_without_a_gpu(tmp_path):
    """The headless host (replaces main.cpp's window + loop) has no CPU fallback either."""
    import subprocess
    import torch
    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present")
    r = subprocess.run([_cli(), "--frames", "1", "--width", "32", "--height", "18", "--out", str(tmp_path / "f")],
                       capture_output=True, text=True)
    assert r.returncode == 1 and "no CPU fallback" in r.stderr
    assert subprocess.run([_cli(), "--bogus"], capture_output=True).returncode == 2


@pytest.mark.gpu
def test_cli_time_sweep_matches_library_and_oracle(tmp_path):
    """rt_cli --builtin sweeps the `time` uniform like the reference's clock (ms-quantised,
    main.cpp:111-118) and writes upright PPMs (rows flipped, alpha dropped, frag:9)."""
    import json
    import subprocess
    import oracle_binding as ob
    from parity import assert_parity
    w, h = 160, 90
    r = subprocess.run([_cli(), "--time", "0.5", "--dt", "0.3337", "--frames", "3", "--width", str(w), "--height", str(h),
                        "--depth", "2", "--out", str(tmp_path / "f"), "--save-scene", str(tmp_path / "s.rtscene")],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    lines = [json.loads(x) for x in r.stdout.splitlines()]
    assert [x["frame"] for x in lines] == [0, 1, 2]
    with rt.Raytracer(0) as tr:
        for k, x in enumerate(lines):
            ms = int(0.5 * 1000.0 + k * float(np.float32(0.3337)) * 1000.0 + 0.5)
            t = float(np.float32(ms) / np.float32(1000.0))
            assert abs(x["time"] - t) < 1e-3 and x["rays"] > 0
            raw = open(x["file"], "rb").read()
            head = b"P6\n%d %d\n255\n" % (w, h)
            assert raw.startswith(head)
            rgb = np.frombuffer(raw[len(head):], np.uint8).reshape(h, w, 3)[::-1]      # PPM is upright; GL order is bottom-up
            sc, cam = rt.builtin_scene(t, 2), rt.builtin_camera(t)
            img = tr.render(sc, cam, w, h)
            assert np.array_equal(rgb, img[..., :3])
            assert_parity(img, ob.render(sc, cam, w, h), "cli frame %d" % k)
        # the saved scene file renders to the same first frame through --scene
        r2 = subprocess.run([_cli(), "--scene", str(tmp_path / "s.rtscene"), "--width", str(w), "--height", str(h),
                             "--format", "raw", "--out", str(tmp_path / "g"), "--quiet"], capture_output=True, text=True)
        assert r2.returncode == 0, r2.stderr
        sc0, cam0 = rt.load_scene(str(tmp_path / "s.rtscene"))
        got = np.fromfile(str(tmp_path / "g_0000.rgba"), np.uint8).reshape(h, w, 4)
        assert np.array_equal(got, tr.render(sc0, cam0, w, h)) and sc0.max_depth == 2
