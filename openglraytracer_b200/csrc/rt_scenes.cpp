/*
 * rt_scenes.cpp — the shader's hard-coded data as host constructors, the scaled
 * synthetic scene, and the PPM/raw dump that replaces the fullscreen-quad blit.
 *
 *   rt_builtin_scene  : glsl:74-157 materials, glsl:199-224 lights, glsl:261-321 objects
 *   rt_builtin_camera : glsl:334-364
 *   rt_write_ppm      : main.cpp:240-260 + draw_screen_frag.glsl:9 (rgb only, alpha forced opaque)
 * Compiled with -ffp-contract=off (time-animated values must match the shader's op order).
 */
#include <cmath>
#include <cstdio>
#include <cstring>

#include "../../include/rt_b200.h"

namespace {

rt_material material(float amb, float dr, float dg, float db, float da, float spec, float shin,
                     float refl, float transp, float ior) {
    rt_material m;
    memset(&m, 0, sizeof(m));
    for (int i = 0; i < 4; i++) { m.ambient[i] = amb; m.specular[i] = spec; }
    m.diffuse[0] = dr; m.diffuse[1] = dg; m.diffuse[2] = db; m.diffuse[3] = da;
    m.shininess = shin; m.reflectivity = refl; m.transparency = transp; m.refraction_index = ior;
    return m;
}
const rt_material kRedGlass = material(1.0f, 1.0f, 0.0f, 0.0f, 1.0f, 1.0f, 10.0f, 0.8f, 0.4f, 1.5f);    // glsl:98-108
const rt_material kGreenGlass = material(1.0f, 0.0f, 1.0f, 0.0f, 1.0f, 1.0f, 10.0f, 0.4f, 0.6f, 1.5f);  // glsl:110-120
const rt_material kBlueGlass = material(1.0f, 0.0f, 0.0f, 1.0f, 1.0f, 1.0f, 10.0f, 0.4f, 0.6f, 1.5f);   // glsl:122-132
const rt_material kMirror = material(1.0f, 0.6f, 0.6f, 0.6f, 1.0f, 1.0f, 4.0f, 1.0f, 0.0f, 1.0f);       // glsl:135-145
const rt_material kWall = material(0.5f, 0.4f, 0.4f, 0.4f, 0.4f, 0.3f, 3.0f, 0.3f, 0.0f, 1.0f);         // glsl:147-157

void box(rt_object* o, float x0, float y0, float z0, float x1, float y1, float z1) {
    o->box_mins[0] = x0; o->box_mins[1] = y0; o->box_mins[2] = z0;
    o->box_maxs[0] = x1; o->box_maxs[1] = y1; o->box_maxs[2] = z1;
    o->sphere_radius = -1.0f;   // null_sphere, glsl:179
}
void place(rt_object* o, float px, float py, float pz, float pitch, float yaw, float roll) {
    o->position[0] = px; o->position[1] = py; o->position[2] = pz;
    o->angles[0] = pitch; o->angles[1] = yaw; o->angles[2] = roll;
}

void builtin_lights(rt_light* L) {
    memset(L, 0, 3 * sizeof(rt_light));
    const float pos[3][3] = {{0.1f, 0.1f, 0.1f}, {7.0f, 7.0f, 2.0f}, {3.0f, -3.0f, 4.0f}};
    const float amb[3] = {0.3f, 0.05f, 0.05f};
    for (int j = 0; j < 3; j++) {
        memcpy(L[j].position, pos[j], 12);
        for (int k = 0; k < 4; k++) L[j].ambient[k] = amb[j];
    }
    for (int k = 0; k < 4; k++) { L[1].diffuse[k] = 1.0f; L[1].specular[k] = 1.0f; }    // glsl:209-214
    L[2].diffuse[0] = 1.0f; L[2].diffuse[3] = 1.0f;                                     // glsl:216-221
    L[2].specular[0] = 1.0f; L[2].specular[3] = 1.0f;
}

struct SplitMix64 {
    uint64_t s;
    uint64_t next() {
        uint64_t z = (s += 0x9E3779B97F4A7C15ull);
        z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
        z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
        return z ^ (z >> 31);
    }
    double unit() { return (double)(next() >> 11) * (1.0 / 9007199254740992.0); }   // [0,1)
    float range(double a, double b) { return (float)(a + (b - a) * unit()); }
};

}  // namespace

extern "C" {

int rt_builtin_scene(float time, rt_object* obj, rt_light* lights, rt_scene* scene) {
    if (!obj || !lights || !scene) return RT_ERR_BAD_ARG;
    const float scaled_time = time * 0.4f;   // glsl:236-237
    memset(obj, 0, 5 * sizeof(rt_object));
    // 0: large cube surrounding the scene (glsl:264-273)
    box(&obj[0], -11.0f, -11.0f, -11.0f, 11.0f, 11.0f, 11.0f);
    obj[0].material = kWall;
    // 1: mirror cube, pulsing, bobbing, spinning (glsl:275-285)
    const float k = 0.5f * sinf(scaled_time * 0.5f) + 1.5f;
    box(&obj[1], -1.0f * k, -1.0f * k, -1.0f * k, 1.0f * k, 1.0f * k, 1.0f * k);
    place(&obj[1], 0.0f, 0.0f, sinf(scaled_time * 3.0f), 0.0f, scaled_time * 90.0f, 0.0f);
    obj[1].material = kMirror;
    // 2: leaning green slab (glsl:287-297)
    box(&obj[2], -10.0f, -10.0f, -1.0f, 10.0f, 10.0f, 1.0f);
    place(&obj[2], 0.0f, 0.0f, -3.0f, sinf(scaled_time * 5.0f) * 10.0f, 45.0f, 0.0f);
    obj[2].material = kGreenGlass;
    // 3: tumbling blue box (glsl:299-309)
    box(&obj[3], -1.0f, -1.0f, -2.0f, 1.0f, 1.0f, 2.0f);
    place(&obj[3], 3.0f, 4.0f, 1.0f, 45.0f + scaled_time * 45.0f, 0.0f, 45.0f + scaled_time * 180.0f);
    obj[3].material = kBlueGlass;
    // 4: red glass sphere (glsl:311-319); box stays null_box
    obj[4].sphere_radius = 2.0f;
    place(&obj[4], -3.0f, 4.0f, 1.0f, 0.0f, 0.0f, 0.0f);
    obj[4].material = kRedGlass;

    builtin_lights(lights);
    scene->objects = obj; scene->objects_count = 5;      // glsl:321
    scene->lights = lights; scene->lights_count = 3;     // glsl:224
    scene->max_depth = 0;                                // glsl:22
    return RT_OK;
}

int rt_orbit_camera(float speed, rt_camera* c) {
    if (!c) return RT_ERR_BAD_ARG;
    const float radius = 10.0f;                          // glsl:339
    c->position[0] = radius * cosf(speed);               // glsl:344
    c->position[1] = radius * sinf(speed);
    c->position[2] = 0.0f;
    const float q = 1.0f * speed * (180.0f / 3.1416f);   // glsl:353 (3.1416, not PI)
    const float yaw = (q - 360.0f * floorf(q / 360.0f)) + 90.0f;   // mod(q, 360) + 90
    c->angles[0] = 0.0f; c->angles[1] = yaw; c->angles[2] = 0.0f;  // glsl:346-358
    c->near_plane = 0.1f; c->far_plane = 1000.0f;        // glsl:361-362
    c->aspect = 16.0f / 9.0f; c->v_fov = 90.0f;          // glsl:363-364
    return RT_OK;
}

int rt_builtin_camera(float time, rt_camera* c) {
    return rt_orbit_camera(time * 0.4f + 0.5f, c);       // glsl:343
}

int rt_scaled_scene(int n, uint64_t seed, rt_object* obj, rt_light* lights, rt_scene* scene) {
    if (n < 1 || !obj || !lights || !scene) return RT_ERR_BAD_ARG;
    memset(obj, 0, (size_t)n * sizeof(rt_object));
    box(&obj[0], -11.0f, -11.0f, -11.0f, 11.0f, 11.0f, 11.0f);
    obj[0].material = kWall;
    const rt_material cycle[5] = {kMirror, kGreenGlass, kBlueGlass, kRedGlass, kWall};
    SplitMix64 rng{seed};
    const double s = 4.0 * pow((double)n, -1.0 / 3.0);
    for (int i = 1; i < n; i++) {
        rt_object* o = &obj[i];
        const float px = rng.range(-9.0, 9.0), py = rng.range(-9.0, 9.0), pz = rng.range(-9.0, 9.0);
        if (i % 5 == 4) {
            o->sphere_radius = rng.range(0.5 * s, 1.5 * s);
            place(o, px, py, pz, 0.0f, 0.0f, 0.0f);
        } else {
            const float hx = rng.range(0.5 * s, 1.5 * s), hy = rng.range(0.5 * s, 1.5 * s), hz = rng.range(0.5 * s, 1.5 * s);
            box(o, -hx, -hy, -hz, hx, hy, hz);
            place(o, px, py, pz, rng.range(0.0, 360.0), rng.range(0.0, 360.0), rng.range(0.0, 360.0));
        }
        o->material = cycle[i % 5];
    }
    builtin_lights(lights);
    scene->objects = obj; scene->objects_count = n;
    scene->lights = lights; scene->lights_count = 3;
    scene->max_depth = 0;
    return RT_OK;
}

int rt_write_ppm(const char* path, const uint8_t* rgba, int width, int height) {
    if (!path || !rgba || width <= 0 || height <= 0) return RT_ERR_BAD_ARG;
    FILE* f = fopen(path, "wb");
    if (!f) return RT_ERR_BAD_ARG;
    fprintf(f, "P6\n%d %d\n255\n", width, height);
    // image row 0 is the bottom scanline (uv (0,0) -> NDC (-1,-1), main.cpp:166-184): flip
    for (int y = height - 1; y >= 0; y--) {
        const uint8_t* row = rgba + (size_t)y * width * 4;
        for (int x = 0; x < width; x++) fwrite(row + 4 * x, 1, 3, f);
    }
    fclose(f);
    return RT_OK;
}

int rt_write_raw(const char* path, const uint8_t* rgba, int width, int height) {
    if (!path || !rgba || width <= 0 || height <= 0) return RT_ERR_BAD_ARG;
    FILE* f = fopen(path, "wb");
    if (!f) return RT_ERR_BAD_ARG;
    size_t n = fwrite(rgba, 1, (size_t)width * height * 4, f);
    fclose(f);
    return n == (size_t)width * height * 4 ? RT_OK : RT_ERR_BAD_ARG;
}

}  // extern "C"
