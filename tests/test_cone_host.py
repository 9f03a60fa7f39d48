"""The packet tests of the large-scene kernels (openglraytracer_b200/csrc/rt_cone.h: cone, reach, fan plane, grown
radii) must be CONSERVATIVE: they may only reject an object that none of the packet's rays can hit -- that is what
keeps the frames byte-identical with the tests on or off.  The header compiles as plain C++, so the property is
checked here on the CPU (no GPU needed) against an exact double-precision segment / sphere test, on packets built
the way the kernels build them.  The GPU side of the same claim is tests/test_gpu_parity.py (identity of whole
frames under every no_cull switch)."""
import os
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _build(tmp_path, source):
    exe = tmp_path / "cone_host"
    subprocess.check_call(["g++", "-O2", "-ffp-contract=off", "-std=c++17", "-I",
                           os.path.join(ROOT, "openglraytracer_b200", "csrc"), str(source), "-o", str(exe)])
    return exe


def test_packet_tests_are_conservative(tmp_path):
    exe = _build(tmp_path, os.path.join(ROOT, "tests", "cone_host.cpp"))
    cases, must, bad, member_pass, group_bad = (int(x) for x in subprocess.check_output([str(exe), "3000"]).split())
    assert cases == 3000 and must > 300000        # objects some ray of the packet really hits
    assert bad == 0                               # ... none of them rejected by cone / reach / fan plane
    # groups of neighbouring objects (rt_groups.cu, rt_cone.h: group_radius): whenever a member passes a cone or
    # ball test, so does the sphere of its group
    assert member_pass > 300000 and group_bad == 0


def test_the_property_test_has_teeth(tmp_path):
    """A fan plane 30 % too thin, or bounding radii 20 % too small, must show up as violations; so must group
    spheres 10 % too small."""
    src = open(os.path.join(ROOT, "tests", "cone_host.cpp")).read()
    for old, new in (("cone_fan(P.k, nrm, hmax, pad1);", "cone_fan(P.k, nrm, hmax * 0.7f, pad1);"),
                     ("const float rp = packet_radius(", "const float rp = 0.8f * packet_radius(")):
        assert src.count(old) == 1
        mutant = tmp_path / "mutant.cpp"
        mutant.write_text(src.replace(old, new))
        exe = _build(tmp_path, mutant)
        _, _, bad, _, _ = (int(x) for x in subprocess.check_output([str(exe), "600"]).split())
        assert bad > 0
    old = "const float rg = group_radius(P.k.apex, grp, P.grow);"
    assert src.count(old) == 1
    mutant = tmp_path / "mutant.cpp"
    mutant.write_text(src.replace(old, "const float rg = 0.9f * group_radius(P.k.apex, grp, P.grow);"))
    exe = _build(tmp_path, mutant)
    _, _, _, _, group_bad = (int(x) for x in subprocess.check_output([str(exe), "600"]).split())
    assert group_bad > 0
