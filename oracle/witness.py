"""witness.py -- a SECOND, independently written restatement of raytrace_compute.glsl.

TEST INFRASTRUCTURE ONLY (tests/ import it; the product never does).

Why it exists: the reference ships no goldens and its shader cannot run on this box
(no GL stack), so oracle/rt_oracle.cpp -- the FP32, line-by-line, stack-machine
restatement the CUDA kernels are held to -- can only be checked against itself:
PARITY UNPINNED.  This file does not change that.  What it removes is the
single-transcription risk: it was written from the shader text alone, in a different
language, precision and structure, and shares no code and no tables with oracle/
rt_oracle.cpp, oracle/glsl_shim.h or openglraytracer_b200/csrc/rt_scenes.cpp:

  * float64 numpy instead of FP32 C++;
  * a RECURSIVE raytrace over whole wavefronts of rays (the shader emulates that
    recursion with its 100-frame Stack_Element machine, glsl:848-1105);
  * row/column math-notation matrices and numpy.linalg.inv instead of the column-major
    adjugate;
  * the scene / material / light / camera constants re-typed from glsl:74-157,
    199-224, 261-321, 334-364.

tests/test_witness.py holds the oracle to it on whole frames and block by block.

Citations: glsl:N = /root/reference/OpenGLRaytracer/raytrace_compute.glsl line N.
"""
import math

import numpy as np

# glsl:13, 17
PI = 3.14159265358
DEG = PI / 180.0


# ----------------------------------------------------------------------------- scene data
class Mat:
    """glsl:56-69"""

    def __init__(self, amb, dif, spe, shin, emi, refl, transp, ior):
        def v4(x):
            return np.array([x] * 4 if np.isscalar(x) else x, float)
        self.ambient, self.diffuse, self.specular, self.emissive = v4(amb), v4(dif), v4(spe), v4(emi)
        self.shininess, self.reflectivity, self.transparency, self.ior = float(shin), float(refl), float(transp), float(ior)


# glsl:98-157 (material1 / material2, glsl:74-96, are never referenced)
RED_GLASS = Mat(1.0, (1.0, 0.0, 0.0, 1.0), 1.0, 10.0, 0.0, 0.8, 0.4, 1.5)
GREEN_GLASS = Mat(1.0, (0.0, 1.0, 0.0, 1.0), 1.0, 10.0, 0.0, 0.4, 0.6, 1.5)
BLUE_GLASS = Mat(1.0, (0.0, 0.0, 1.0, 1.0), 1.0, 10.0, 0.0, 0.4, 0.6, 1.5)
MIRROR = Mat(1.0, (0.6, 0.6, 0.6, 1.0), 1.0, 4.0, 0.0, 1.0, 0.0, 1.0)
WALL = Mat(0.5, 0.4, 0.3, 3.0, 0.0, 0.3, 0.0, 1.0)


class Obj:
    """glsl:244-258; box = (mins, maxs) or None, sphere = radius or None (glsl:178-179 nulls)"""

    def __init__(self, box, sphere, position, angles, material):
        self.box = None if box is None else (np.array(box[0], float), np.array(box[1], float))
        self.sphere = sphere
        self.position = np.array(position, float)
        self.angles = np.array(angles, float)      # pitch, yaw, roll (degrees)
        self.material = material


class Lamp:
    """glsl:190-196"""

    def __init__(self, position, amb, dif, spe):
        def v4(x):
            return np.array([x] * 4 if np.isscalar(x) else x, float)
        self.position = np.array(position, float)
        self.ambient, self.diffuse, self.specular = v4(amb), v4(dif), v4(spe)


def builtin_lights():
    """glsl:199-222"""
    return [Lamp((0.1, 0.1, 0.1), 0.3, 0.0, 0.0),
            Lamp((7.0, 7.0, 2.0), 0.05, 1.0, 1.0),
            Lamp((3.0, -3.0, 4.0), 0.05, (1.0, 0.0, 0.0, 1.0), (1.0, 0.0, 0.0, 1.0))]


def builtin_objects(time):
    """glsl:236-237, 261-320: the five objects at uniform `time`"""
    st = float(time) * 0.4
    grow = 0.5 * math.sin(st * 0.5) + 1.5
    return [
        Obj(((-11.0,) * 3, (11.0,) * 3), None, (0.0, 0.0, 0.0), (0.0, 0.0, 0.0), WALL),
        Obj(((-grow,) * 3, (grow,) * 3), None, (0.0, 0.0, math.sin(st * 3.0)), (0.0, st * 90.0, 0.0), MIRROR),
        Obj(((-10.0, -10.0, -1.0), (10.0, 10.0, 1.0)), None, (0.0, 0.0, -3.0),
            (math.sin(st * 5.0) * 10.0, 45.0, 0.0), GREEN_GLASS),
        Obj(((-1.0, -1.0, -2.0), (1.0, 1.0, 2.0)), None, (3.0, 4.0, 1.0),
            (45.0 + st * 45.0, 0.0, 45.0 + st * 180.0), BLUE_GLASS),
        Obj(None, 2.0, (-3.0, 4.0, 1.0), (0.0, 0.0, 0.0), RED_GLASS),
    ]


class Cam:
    """glsl:36-50"""

    def __init__(self, position, angles, v_fov, aspect, near, far):
        self.position = np.array(position, float)
        self.angles = np.array(angles, float)
        self.v_fov, self.aspect, self.near, self.far = float(v_fov), float(aspect), float(near), float(far)


def builtin_camera(time):
    """glsl:334-364: orbit of radius 10, looking at the centre; note 3.1416, not PI (glsl:353)"""
    speed = float(time) * 0.4 + 0.5
    yaw = math.fmod(speed * (180.0 / 3.1416), 360.0) + 90.0     # speed >= 0 here: fmod == GLSL mod
    if speed < 0.0:
        x = speed * (180.0 / 3.1416)
        yaw = x - 360.0 * math.floor(x / 360.0) + 90.0
    return Cam((10.0 * math.cos(speed), 10.0 * math.sin(speed), 0.0), (0.0, yaw, 0.0), 90.0, 16.0 / 9.0, 0.1, 1000.0)


def from_structs(scene, camera=None):
    """Scene data handed over as the C-ABI structs (include/rt_b200.h layout; ctypes objects of the
    test bindings): used for scenes that are DATA, not shader text (the scaled scene, edge cases)."""
    objs = []
    for i in range(scene.objects_count):
        o = scene.objects[i]
        m = o.material
        mat = Mat(list(m.ambient), list(m.diffuse), list(m.specular), m.shininess, list(m.emissive),
                  m.reflectivity, m.transparency, m.refraction_index)
        mins, maxs = list(o.box_mins), list(o.box_maxs)
        is_box = any(v != 0.0 for v in mins + maxs)                 # glsl:749: box != null_box
        is_sph = (not is_box) and o.sphere_radius != -1.0           # glsl:759
        objs.append(Obj((mins, maxs) if is_box else None, float(o.sphere_radius) if is_sph else None,
                        list(o.position), list(o.angles), mat))
    lamps = [Lamp(list(scene.lights[j].position), list(scene.lights[j].ambient), list(scene.lights[j].diffuse),
                  list(scene.lights[j].specular)) for j in range(scene.lights_count)]
    cam = None
    if camera is not None:
        cam = Cam(list(camera.position), list(camera.angles), camera.v_fov, camera.aspect, camera.near_plane,
                  camera.far_plane)
    return objs, lamps, cam


# ----------------------------------------------------------------------------- matrices (math notation M[row, col])
def rot_x(deg):
    c, s = math.cos(DEG * deg), math.sin(DEG * deg)      # glsl:444-454
    return np.array([[1, 0, 0, 0], [0, c, -s, 0], [0, s, c, 0], [0, 0, 0, 1.0]])


def rot_y(deg):
    c, s = math.cos(DEG * deg), math.sin(DEG * deg)      # glsl:460-470
    return np.array([[c, 0, s, 0], [0, 1, 0, 0], [-s, 0, c, 0], [0, 0, 0, 1.0]])


def rot_z(deg):
    c, s = math.cos(DEG * deg), math.sin(DEG * deg)      # glsl:476-486
    return np.array([[c, -s, 0, 0], [s, c, 0, 0], [0, 0, 1, 0], [0, 0, 0, 1.0]])


def transform(position, angles):
    """glsl:492-503, 529-532: T * Rz(yaw) * Rx(pitch) * Ry(roll)"""
    t = np.eye(4)
    t[:3, 3] = position
    return t @ rot_z(angles[1]) @ rot_x(angles[0]) @ rot_y(angles[2])


def projection(cam):
    """glsl:411-426 (GLSL result[col][row]: [2][3] = -1 is row 3, col 2; [3][2] = C is row 2, col 3)"""
    q = 1.0 / math.tan(DEG * 0.5 * cam.v_fov)
    m = np.zeros((4, 4))
    m[0, 0] = q / cam.aspect
    m[1, 1] = q
    m[2, 2] = (cam.near + cam.far) / (cam.near - cam.far)
    m[3, 2] = -1.0
    m[2, 3] = (2.0 * cam.near * cam.far) / (cam.near - cam.far)
    return m


def view(cam):
    """glsl:538-545"""
    return np.linalg.inv(transform(cam.position, cam.angles) @ rot_x(90.0))


def primary_rays(cam, width, height, xs, ys):
    """glsl:377-392 for pixel coordinate arrays xs, ys -> (origins, directions)"""
    hw, hh = width // 2, height // 2                      # integer halves, no pixel-centre offset
    vx = (xs - hw).astype(float) / float(hw) if hw else np.full(xs.shape, np.nan)
    vy = (ys - hh).astype(float) / float(hh) if hh else np.full(ys.shape, np.nan)
    if hw == 0:                                           # float(x - 0) / float(0): +-inf or nan like the shader
        with np.errstate(all="ignore"):
            vx = (xs - hw).astype(float) / np.zeros(xs.shape)
    if hh == 0:
        with np.errstate(all="ignore"):
            vy = (ys - hh).astype(float) / np.zeros(ys.shape)
    unproject = np.linalg.inv(projection(cam) @ view(cam))
    n = xs.size
    a = np.stack([vx, vy, np.full(n, 0.5), np.ones(n)], axis=1) @ unproject.T
    b = np.stack([vx, vy, np.ones(n), np.ones(n)], axis=1) @ unproject.T
    with np.errstate(all="ignore"):
        d = (b / b[:, 3:4] - a / a[:, 3:4])[:, :3]
        d = d / np.sqrt((d * d).sum(axis=1, keepdims=True))
    o = np.broadcast_to(cam.position, (n, 3)).copy()
    return o, d


# ----------------------------------------------------------------------------- intersections
def _dot(a, b):
    return (a * b).sum(axis=1)


def hit_sphere(o, d, obj):
    """glsl:583-640 for ray arrays; returns t (<= 0 or nan: nothing usable), p, n, inside"""
    with np.errstate(all="ignore"):
        oc = o - obj.position
        qa = _dot(d, d)
        qb = _dot(2.0 * d, oc)
        qc = _dot(oc, oc) - obj.sphere * obj.sphere
        qd = qb * qb - 4.0 * qa * qc
        root = np.sqrt(np.where(qd < 0.0, 0.0, qd))
        t1 = (-qb + root) / (2.0 * qa)
        t2 = (-qb - root) / (2.0 * qa)
        lo, hi = np.fmin(t1, t2), np.fmax(t1, t2)
        inside = lo < 0.0
        t = np.where(inside, hi, lo)
        t = np.where((qd < 0.0) | (hi < 0.0), -1.0, t)
        p = o + t[:, None] * d
        n = p - obj.position
        n = n / np.sqrt(_dot(n, n))[:, None]
        n = np.where(inside[:, None], -n, n)
    return t, p, n, inside & (t > 0)


def hit_box(o, d, obj):
    """glsl:647-724"""
    l2w = transform(obj.position, obj.angles)
    w2l = np.linalg.inv(l2w)
    mins, maxs = obj.box
    with np.errstate(all="ignore"):
        s = o @ w2l[:3, :3].T + w2l[:3, 3]
        e = d @ w2l[:3, :3].T
        ta = (mins - s) / e
        tb = (maxs - s) / e
        t1, t2 = np.fmin(ta, tb), np.fmax(ta, tb)
        near = np.fmax(np.fmax(t1[:, 0], t1[:, 1]), t1[:, 2])
        far = np.fmin(np.fmin(t2[:, 0], t2[:, 1]), t2[:, 2])
        miss = (near >= far) | (far <= 0.0)
        inside = near < 0.0
        t = np.where(inside, far, near)
        bound = np.where(inside[:, None], t2, t1)
        face = np.where(t == bound[:, 1], 1, np.where(t == bound[:, 2], 2, 0))       # y wins, then z, else x
        rows = np.arange(o.shape[0])
        nl = np.zeros_like(o)
        nl[rows, face] = 1.0
        nl = np.where((e[rows, face] > 0.0)[:, None], -nl, nl)
        nmat = np.linalg.inv(l2w[:3, :3]).T                                             # glsl:718, not renormalised
        n = nl @ nmat.T
        q = s + t[:, None] * e
        p = q @ l2w[:3, :3].T + l2w[:3, 3]                                              # glsl:721
        t = np.where(miss, -1.0, t)
    return t, p, n, inside & ~miss


def closest(o, d, objs):
    """glsl:738-782: index (-1 = none), t, p, n, inside per ray; strict '<' so the first index wins a tie"""
    n_rays = o.shape[0]
    best = np.full(n_rays, 10000.0)
    idx = np.full(n_rays, -1)
    bp, bn = np.zeros((n_rays, 3)), np.zeros((n_rays, 3))
    bin_ = np.zeros(n_rays, bool)
    for i, obj in enumerate(objs):
        if obj.box is not None:
            t, p, n, ins = hit_box(o, d, obj)
        elif obj.sphere is not None:
            t, p, n, ins = hit_sphere(o, d, obj)
        else:
            continue
        with np.errstate(invalid="ignore"):
            take = ~(t <= 0.0) & (t < best)
        best = np.where(take, t, best)
        idx = np.where(take, i, idx)
        bp = np.where(take[:, None], p, bp)
        bn = np.where(take[:, None], n, bn)
        bin_ = np.where(take, ins, bin_)
    return idx, best, bp, bn, bin_


# ----------------------------------------------------------------------------- shading and recursion
def reflect(i, n):
    """GLSL reflect: I - 2 dot(N, I) N"""
    return i - 2.0 * _dot(n, i)[:, None] * n


def refract(i, n, eta):
    """GLSL refract: k = 1 - eta^2 (1 - dot(N, I)^2); k < 0 ? vec3(0) : eta I - (eta dot(N, I) + sqrt(k)) N"""
    with np.errstate(all="ignore"):
        ni = _dot(n, i)
        k = 1.0 - eta * eta * (1.0 - ni * ni)
        out = eta[:, None] * i - (eta * ni + np.sqrt(np.where(k < 0.0, 0.0, k)))[:, None] * n
    return np.where((k < 0.0)[:, None], 0.0, out)


def _gather(objs, idx, attr):
    table = np.array([np.atleast_1d(getattr(o.material, attr)) for o in objs], float)
    return table[idx]


def phong(o, d, idx, p, n, objs, lamps, counters):
    """glsl:789-840 for the rays that hit (arrays already restricted to them)"""
    m_amb, m_dif, m_spe, m_emi = (_gather(objs, idx, a) for a in ("ambient", "diffuse", "specular", "emissive"))
    shin = _gather(objs, idx, "shininess")[:, 0]
    amb, dif, spe = np.zeros_like(m_amb), np.zeros_like(m_amb), np.zeros_like(m_amb)
    with np.errstate(all="ignore"):
        vdir = -d / np.sqrt(_dot(d, d))[:, None]
        for lamp in lamps:
            amb += lamp.ambient * m_amb
            to_l = lamp.position - p
            ldir = to_l / np.sqrt(_dot(to_l, to_l))[:, None]
            s_idx, s_t, _, _, _ = closest(p + n * 0.01, to_l, objs)           # un-normalised direction (glsl:809)
            counters["rays"] += o.shape[0]
            lit = ~((s_idx != -1) & (s_t < 1.0))
            refl = reflect(-ldir, n)
            refl = refl / np.sqrt(_dot(refl, refl))[:, None]
            kd = np.maximum(_dot(ldir, n), 0.0)
            ks = np.power(np.maximum(_dot(vdir, refl), 0.0), shin)
            dif += np.where(lit[:, None], lamp.diffuse * m_dif * kd[:, None], 0.0)
            spe += np.where(lit[:, None], lamp.specular * m_spe * ks[:, None], 0.0)
    total = amb + dif + spe + m_emi
    return total[:, :3] * total[:, 3:4]                                        # rgb scaled by the summed alpha (glsl:839)


def raytrace(o, d, depth, objs, lamps, counters):
    """The recursion that glsl:848-1105 emulates with its stack machine.  Returns (colour, loop
    iterations the stack machine would spend: 6 per hit node, 1 per miss node) per ray."""
    n_rays = o.shape[0]
    col = np.zeros((n_rays, 3))
    its = np.ones(n_rays, np.int64)
    if n_rays == 0:
        return col, its
    idx, t, p, n, inside = closest(o, d, objs)
    counters["rays"] += n_rays
    counters["nodes"] += n_rays
    h = np.nonzero(idx >= 0)[0]
    if h.size == 0:
        return col, its
    counters["hit_nodes"] += h.size
    its[h] = 6
    ho, hd, hi, hp, hn, hin = o[h], d[h], idx[h], p[h], n[h], inside[h]
    c = phong(ho, hd, hi, hp, hn, objs, lamps, counters)
    if depth > 0:
        kr = _gather(objs, hi, "reflectivity")[:, 0]
        kt = _gather(objs, hi, "transparency")[:, 0]
        ior = _gather(objs, hi, "ior")[:, 0]
        r = np.nonzero(kr > 0.0)[0]                                             # glsl:979-997
        f = np.nonzero(kt > 0.0)[0]                                             # glsl:1002-1032
        with np.errstate(all="ignore"):
            eta = 1.0 / ior[f]                                                  # air index 1.0 (glsl:72) / object index
            eta = np.where(hin[f], 1.0 / eta, eta)                              # leaving the object: inverted (glsl:1017-1020)
        # both kinds of children of this wavefront go down in ONE recursive call (reflected first)
        co = np.concatenate([hp[r] + hn[r] * 0.001, hp[f] - hn[f] * 0.001])
        cd = np.concatenate([reflect(hd[r], hn[r]), refract(hd[f], hn[f], eta)])   # total internal reflection: vec3(0)
        cc, ci = raytrace(co, cd, depth - 1, objs, lamps, counters)
        # final = mix(mix(phong, reflected, reflectivity), refracted, transparency)  (glsl:1046-1051)
        c[r] = c[r] * (1.0 - kr[r])[:, None] + cc[:r.size] * kr[r][:, None]
        c[f] = c[f] * (1.0 - kt[f])[:, None] + cc[r.size:] * kt[f][:, None]
        np.add.at(its, h[r], ci[:r.size])
        np.add.at(its, h[f], ci[r.size:])
    col[h] = c
    return col, its


def to_rgba8(col):
    """imageStore into an RGBA8 image (glsl:404, main.cpp:223): clamp, *255, round half to even; alpha 0"""
    with np.errstate(invalid="ignore"):
        c = np.where(np.isnan(col), 0.0, np.clip(col, 0.0, 1.0))
    out = np.zeros(col.shape[:-1] + (4,), np.uint8)
    out[..., :3] = np.rint(c * 255.0).astype(np.uint8)
    return out


def render(objs, lamps, cam, width, height, depth, pixels=None, want_float=False, want_stats=False):
    """main() (glsl:325-405) for the whole frame (row 0 = bottom scanline like the image2D) or for a
    list of (x, y) pixels."""
    if pixels is None:
        ys, xs = np.mgrid[0:height, 0:width]
        xs, ys = xs.ravel(), ys.ravel()
    else:
        pix = np.asarray(pixels).reshape(-1, 2)
        xs, ys = pix[:, 0], pix[:, 1]
    counters = {"rays": 0, "nodes": 0, "hit_nodes": 0}
    o, d = primary_rays(cam, width, height, xs, ys)
    col, its = raytrace(o, d, depth, objs, lamps, counters)
    red = its >= 10001                                                         # glsl:1076-1101
    col[red] = (1.0, 0.0, 0.0)
    counters["red_pixels"] = int(red.sum())
    img = to_rgba8(col)
    if pixels is None:
        img = img.reshape(height, width, 4)
        col = col.reshape(height, width, 3)
    res = [img]
    if want_float:
        res.append(col)
    if want_stats:
        res.append(counters)
    return res[0] if len(res) == 1 else tuple(res)


def render_builtin(time, width, height, depth, **kw):
    return render(builtin_objects(time), builtin_lights(), builtin_camera(time), width, height, depth, **kw)
