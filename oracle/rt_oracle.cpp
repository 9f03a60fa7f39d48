/*
 * rt_oracle.cpp — CPU oracle: FP32 restatement of the reference compute shader.
 *
 * TEST INFRASTRUCTURE ONLY (see rt_oracle.h).  PARITY UNPINNED (no reference
 * goldens exist; the shader cannot be executed on this box).
 *
 * Follows /root/reference/OpenGLRaytracer/raytrace_compute.glsl function by
 * function, keeping its redundancy and quirks: the 100-frame Stack_Element
 * machine with one phase per loop iteration (glsl:848-1105), closest-hit shadow
 * rays with an un-normalised direction (glsl:807-819), rgb scaled by the summed
 * alpha (glsl:839), per-test rebuild of the box transform and its inverse
 * (glsl:650-652), per-pixel rebuild of the camera matrices (glsl:366-383).
 * Each function cites the glsl lines it restates.
 *
 * Build: g++ -O2 -ffp-contract=off -fopenmp -shared -fPIC (oracle/Makefile).
 */
#include "rt_oracle.h"
#include "glsl_shim.h"

#include <cstring>
#include <vector>
#ifdef _OPENMP
#include <omp.h>
#endif

using namespace glsl;

namespace {

// glsl:13-17
const float PI = 3.14159265358f;
const float DEG_TO_RAD = PI / 180.0f;
// glsl:72
const float air_index_of_refraction = 1.0f;

// glsl:25-31
struct Ray { vec3 start; vec3 dir; };

// glsl:561-573
struct Collision {
    float t; vec3 p; vec3 n; bool inside; int object_index;
};
// glsl:576
inline Collision null_collision() { Collision c; c.t = -1.0f; c.p = vec3(0.0f); c.n = vec3(0.0f); c.inside = false; c.object_index = -1; return c; }

// glsl:848-864
struct Stack_Element {
    int type; int depth; int counter;
    vec3 phong_color, reflected_color, refracted_color;
    bool reflected, refracted;
    vec3 final_color;
    Ray ray;
    Collision collision;
};
const int RAY_TYPE_REFLECTION = 1;  // glsl:867
const int RAY_TYPE_REFRACTION = 2;  // glsl:868
const int stack_size = 100;         // glsl:874

// glsl:871
inline Stack_Element null_stack_element() {
    Stack_Element e;
    e.type = 0; e.depth = -1; e.counter = -1;
    e.phong_color = vec3(0.0f); e.reflected_color = vec3(0.0f); e.refracted_color = vec3(0.0f);
    e.reflected = false; e.refracted = false; e.final_color = vec3(0.0f);
    e.ray.start = vec3(0.0f); e.ray.dir = vec3(0.0f);
    e.collision = null_collision();
    return e;
}
// GLSL struct `!=` compares every member (glsl:934)
inline bool same(const Collision& a, const Collision& b) {
    return a.t == b.t && a.p == b.p && a.n == b.n && a.inside == b.inside && a.object_index == b.object_index;
}
inline bool same(const Stack_Element& a, const Stack_Element& b) {
    return a.type == b.type && a.depth == b.depth && a.counter == b.counter &&
           a.phong_color == b.phong_color && a.reflected_color == b.reflected_color &&
           a.refracted_color == b.refracted_color && a.reflected == b.reflected &&
           a.refracted == b.refracted && a.final_color == b.final_color &&
           a.ray.start == b.ray.start && a.ray.dir == b.ray.dir && same(a.collision, b.collision);
}

inline vec3 v3(const float* f) { return vec3(f[0], f[1], f[2]); }
inline vec4 v4(const float* f) { return vec4(f[0], f[1], f[2], f[3]); }

// glsl:411-426
mat4 calc_projection_matrix(const rt_camera& c) {
    float q = 1.0f / tanf(DEG_TO_RAD * 0.5f * c.v_fov);
    float A = q / c.aspect;
    float B = (c.near_plane + c.far_plane) / (c.near_plane - c.far_plane);
    float C = (2.0f * c.near_plane * c.far_plane) / (c.near_plane - c.far_plane);
    mat4 result(0.0f);
    result.c[0][0] = A;
    result.c[1][1] = q;
    result.c[2][2] = B;
    result.c[2][3] = -1.0f;
    result.c[3][2] = C;
    return result;
}
// glsl:432-437
mat4 translation_matrix(vec3 t) {
    mat4 r(1.0f);
    r.c[3][0] = t.x; r.c[3][1] = t.y; r.c[3][2] = t.z; r.c[3][3] = 1.0f;
    return r;
}
// glsl:444-454
mat4 rotation_matrix_x(float deg) {
    float cos_deg = cosf(DEG_TO_RAD * deg);
    float sin_deg = sinf(DEG_TO_RAD * deg);
    mat4 r(1.0f);
    r.c[1][1] = cos_deg; r.c[1][2] = sin_deg; r.c[2][1] = -sin_deg; r.c[2][2] = cos_deg;
    return r;
}
// glsl:460-470
mat4 rotation_matrix_y(float deg) {
    float cos_deg = cosf(DEG_TO_RAD * deg);
    float sin_deg = sinf(DEG_TO_RAD * deg);
    mat4 r(1.0f);
    r.c[0][0] = cos_deg; r.c[0][2] = -sin_deg; r.c[2][0] = sin_deg; r.c[2][2] = cos_deg;
    return r;
}
// glsl:476-486
mat4 rotation_matrix_z(float deg) {
    float cos_deg = cosf(DEG_TO_RAD * deg);
    float sin_deg = sinf(DEG_TO_RAD * deg);
    mat4 r(1.0f);
    r.c[0][0] = cos_deg; r.c[0][1] = sin_deg; r.c[1][0] = -sin_deg; r.c[1][1] = cos_deg;
    return r;
}
// glsl:492-503
mat4 rotation_matrix(vec3 r) {
    mat4 result(1.0f);
    result = result * rotation_matrix_z(r.y);
    result = result * rotation_matrix_x(r.x);
    result = result * rotation_matrix_y(r.z);
    return result;
}
// glsl:529-532
mat4 calc_transform_matrix(vec3 position, vec3 angles) {
    return translation_matrix(position) * rotation_matrix(angles);
}
// glsl:538-545
mat4 calc_view_matrix(const rt_camera& c) {
    mat4 flip_y_and_z = rotation_matrix_x(90.0f);
    return inverse(calc_transform_matrix(v3(c.position), v3(c.angles)) * flip_y_and_z);
}

// The pure per-object values intersect_box_object recomputes on every call
// (glsl:650-652, 718).  as_written=1 recomputes them; as_written=0 looks them up.
struct BoxXform { mat4 local_to_world; mat4 world_to_local; mat3 normal_mat; };
inline BoxXform box_xform(const rt_object& o) {
    BoxXform x;
    x.local_to_world = calc_transform_matrix(v3(o.position), v3(o.angles));
    x.world_to_local = inverse(x.local_to_world);
    x.normal_mat = transpose(inverse(mat3(x.local_to_world)));
    return x;
}

// One shader invocation's global state (glsl:873-880 stack, glsl:261-321 scene).
struct Invocation {
    const rt_object* objects; int objects_count;
    const rt_light* lights; int lights_count;
    const BoxXform* cache;   // null in as_written mode
    Stack_Element stack[stack_size];
    int stack_pointer;
    Stack_Element popped_stack_element;
    oracle_stats st;

    Collision intersect_sphere_object(const Ray& r, const rt_object& o);
    Collision intersect_box_object(const Ray& r, const rt_object& o, int index);
    Collision get_closest_collision(const Ray& r);
    vec3 ads_phong_lighting(const Ray& r, const Collision& c);
    void push_stack_element(const Ray& r, int depth, int type);
    void pop_stack_element();
    void process_stack_element(int index);
    vec3 recursive_raytrace(const Ray& r, int max_depth);
};

// glsl:583-640
Collision Invocation::intersect_sphere_object(const Ray& r, const rt_object& o) {
    st.sphere_tests++;
    vec3 opos = v3(o.position);
    float radius = o.sphere_radius;
    float qa = dot(r.dir, r.dir);
    float qb = dot(2.0f * r.dir, r.start - opos);
    float qc = dot(r.start - opos, r.start - opos) - radius * radius;

    float qd = qb * qb - 4.0f * qa * qc;

    Collision c;
    c.t = 0.0f; c.p = vec3(0.0f); c.n = vec3(0.0f); c.object_index = -1;
    c.inside = false;

    if (qd < 0.0f) { c.t = -1.0f; return c; }

    float sqrt_qd = sqrtf(qd);
    float t1 = (-qb + sqrt_qd) / (2.0f * qa);
    float t2 = (-qb - sqrt_qd) / (2.0f * qa);
    float t_near = gmin(t1, t2);
    float t_far = gmax(t1, t2);

    c.t = t_near;
    if (t_far < 0.0f) { c.t = -1.0f; return c; }
    if (t_near < 0.0f) { c.t = t_far; c.inside = true; }

    c.p = r.start + c.t * r.dir;
    c.n = normalize(c.p - opos);
    if (c.inside) c.n = c.n * -1.0f;
    return c;
}

// glsl:647-724
Collision Invocation::intersect_box_object(const Ray& r, const rt_object& o, int index) {
    st.box_tests++;
    BoxXform fresh;
    const BoxXform* x;
    if (cache) {
        x = &cache[index];
    } else {
        // as written: rebuild the transform and invert it for every (ray, box) pair
        fresh.local_to_world = calc_transform_matrix(v3(o.position), v3(o.angles));
        fresh.world_to_local = inverse(fresh.local_to_world);
        x = &fresh;
    }
    const mat4& local_to_world = x->local_to_world;
    const mat4& world_to_local = x->world_to_local;

    vec3 ray_start = (world_to_local * vec4(r.start, 1.0f)).xyz();
    vec3 ray_dir = (world_to_local * vec4(r.dir, 0.0f)).xyz();

    vec3 mins = v3(o.box_mins), maxs = v3(o.box_maxs);
    vec3 t_min = (mins - ray_start) / ray_dir;
    vec3 t_max = (maxs - ray_start) / ray_dir;
    vec3 t1 = gmin(t_min, t_max);
    vec3 t2 = gmax(t_min, t_max);
    float t_near = gmax(gmax(t1.x, t1.y), t1.z);
    float t_far = gmin(gmin(t2.x, t2.y), t2.z);

    Collision c;
    c.p = vec3(0.0f); c.n = vec3(0.0f); c.object_index = -1;
    c.t = t_near;
    c.inside = false;

    if (t_near >= t_far || t_far <= 0.0f) { c.t = -1.0f; return c; }

    float intersection = t_near;
    vec3 boundary = t1;
    if (t_near < 0.0f) {
        c.t = t_far;
        intersection = t_far;
        boundary = t2;
        c.inside = true;
    }

    int face_index = 0;
    if (intersection == boundary.y) face_index = 1;
    else if (intersection == boundary.z) face_index = 2;

    c.n = vec3(0.0f);
    c.n[face_index] = 1.0f;
    if (ray_dir[face_index] > 0.0f) c.n = c.n * -1.0f;

    // glsl:718 (recomputed per hit as written)
    if (cache) c.n = x->normal_mat * c.n;
    else c.n = transpose(inverse(mat3(local_to_world))) * c.n;

    c.p = (local_to_world * vec4(ray_start + c.t * ray_dir, 1.0f)).xyz();
    return c;
}

// glsl:738-782
Collision Invocation::get_closest_collision(const Ray& r) {
    st.rays++;
    float closest = 10000.0f;
    Collision closest_collision = null_collision();
    closest_collision.object_index = -1;

    for (int i = 0; i < objects_count; i++) {
        const rt_object& o = objects[i];
        Collision c;
        bool is_box = !(o.box_mins[0] == 0.0f && o.box_mins[1] == 0.0f && o.box_mins[2] == 0.0f &&
                        o.box_maxs[0] == 0.0f && o.box_maxs[1] == 0.0f && o.box_maxs[2] == 0.0f);  // glsl:749
        if (is_box) {
            c = intersect_box_object(r, o, i);
            if (c.t <= 0.0f) continue;
        } else if (o.sphere_radius != -1.0f) {  // glsl:759
            c = intersect_sphere_object(r, o);
            if (c.t <= 0.0f) continue;
        } else {
            continue;
        }
        if (c.t < closest) {
            closest = c.t;
            closest_collision = c;
            closest_collision.object_index = i;
        }
    }
    return closest_collision;
}

// glsl:789-840
vec3 Invocation::ads_phong_lighting(const Ray& r, const Collision& c) {
    const rt_material& mat = objects[c.object_index].material;
    vec4 ambient(0.0f), diffuse(0.0f), specular(0.0f);

    for (int j = 0; j < lights_count; j++) {
        const rt_light& L = lights[j];
        ambient = ambient + v4(L.ambient) * v4(mat.ambient);

        vec3 light_dir = normalize(v3(L.position) - c.p);

        Ray light_ray;
        light_ray.start = c.p + c.n * 0.01f;
        light_ray.dir = v3(L.position) - c.p;
        bool in_shadow = false;

        st.shadow_rays++;
        Collision c_shadow = get_closest_collision(light_ray);
        if (c_shadow.object_index != -1 && c_shadow.t < 1.0f) in_shadow = true;

        if (in_shadow == false) {
            vec3 light_ref = normalize(reflect(-light_dir, c.n));
            float cos_theta = dot(light_dir, c.n);
            float cos_phi = dot(normalize(-r.dir), light_ref);
            diffuse = diffuse + v4(L.diffuse) * v4(mat.diffuse) * gmax(cos_theta, 0.0f);
            specular = specular + v4(L.specular) * v4(mat.specular) * powf(gmax(cos_phi, 0.0f), mat.shininess);
        }
    }
    vec4 phong_color = ambient + diffuse + specular + v4(mat.emissive);
    return phong_color.rgb() * phong_color.w;
}

// glsl:886-909
void Invocation::push_stack_element(const Ray& r, int depth, int type) {
    if (stack_pointer >= stack_size) return;
    Stack_Element element;
    element.type = type;
    element.depth = depth;
    element.counter = 0;
    element.phong_color = vec3(0.0f);
    element.reflected_color = vec3(0.0f);
    element.refracted_color = vec3(0.0f);
    element.reflected = false;
    element.refracted = false;
    element.final_color = vec3(0.0f);
    element.ray = r;
    element.collision = null_collision();
    stack[stack_pointer] = element;
    stack_pointer++;
    st.nodes++;
}

// glsl:914-922
void Invocation::pop_stack_element() {
    stack_pointer--;
    popped_stack_element = stack[stack_pointer];
    stack[stack_pointer] = null_stack_element();
}

// glsl:930-1065
void Invocation::process_stack_element(int index) {
    if (!same(popped_stack_element, null_stack_element())) {
        if (popped_stack_element.type == RAY_TYPE_REFLECTION)
            stack[index].reflected_color = popped_stack_element.final_color;
        else if (popped_stack_element.type == RAY_TYPE_REFRACTION)
            stack[index].refracted_color = popped_stack_element.final_color;
        popped_stack_element = null_stack_element();
    }

    Ray r = stack[index].ray;
    Collision c = stack[index].collision;

    while (stack[index].counter < 5) {
        // PHASE 0 (glsl:956-967)
        if (stack[index].counter == 0) {
            c = get_closest_collision(r);
            if (c.object_index == -1) break;
            stack[index].collision = c;
            st.hit_nodes++;
        }
        // PHASE 1 (glsl:972-975)
        if (stack[index].counter == 1) {
            stack[index].phong_color = ads_phong_lighting(r, c);
        }
        // PHASE 2 (glsl:979-997)
        if (stack[index].counter == 2) {
            if (stack[index].depth > 0) {
                if (objects[c.object_index].material.reflectivity > 0.0f) {
                    Ray reflected_ray;
                    reflected_ray.start = c.p + c.n * 0.001f;
                    reflected_ray.dir = reflect(r.dir, c.n);
                    push_stack_element(reflected_ray, stack[index].depth - 1, RAY_TYPE_REFLECTION);
                    stack[index].reflected = true;
                }
            }
        }
        // PHASE 3 (glsl:1001-1030)
        if (stack[index].counter == 3) {
            if (stack[index].depth > 0) {
                if (objects[c.object_index].material.transparency > 0.0f) {
                    Ray refracted_ray;
                    refracted_ray.start = c.p - c.n * 0.001f;
                    float object_index_of_refraction = objects[c.object_index].material.refraction_index;
                    float refraction_ratio = air_index_of_refraction / object_index_of_refraction;
                    if (c.inside) refraction_ratio = 1.0f / refraction_ratio;
                    refracted_ray.dir = refract(r.dir, c.n, refraction_ratio);
                    push_stack_element(refracted_ray, stack[index].depth - 1, RAY_TYPE_REFRACTION);
                    stack[index].refracted = true;
                }
            }
        }
        // PHASE 4 (glsl:1034-1054)
        if (stack[index].counter == 4) {
            const rt_material& mat = objects[c.object_index].material;
            vec3 phong_color = stack[index].phong_color;
            vec3 reflected_color = stack[index].reflected_color;
            vec3 refracted_color = stack[index].refracted_color;
            vec3 final_color = phong_color;
            if (stack[index].reflected) final_color = mix(final_color, reflected_color, mat.reflectivity);
            if (stack[index].refracted) final_color = mix(final_color, refracted_color, mat.transparency);
            stack[index].final_color = final_color;
        }
        stack[index].counter++;
        return;  // one phase per invocation (glsl:1059-1060)
    }
    pop_stack_element();
}

// glsl:1071-1105
vec3 Invocation::recursive_raytrace(const Ray& r, int max_depth) {
    // Per-invocation initial state of the shader globals (glsl:877, 880):
    // stack_pointer = 0; popped_stack_element has no initialiser -> zero.
    stack_pointer = 0;
    std::memset((void*)&popped_stack_element, 0, sizeof(popped_stack_element));

    push_stack_element(r, max_depth, RAY_TYPE_REFLECTION);
    int break_counter = 0;
    int break_at = 10000;
    while (stack_pointer > 0) {
        if (break_counter++ > break_at) break;
        int element_index = stack_pointer - 1;
        st.loop_iterations++;
        process_stack_element(element_index);
    }
    if (break_counter > break_at) { st.red_pixels++; return vec3(1.0f, 0.0f, 0.0f); }
    return popped_stack_element.final_color;
}

// Jitter extension (NOT in the reference, which samples the
// integer pixel corner only, glsl:377-378).  Exact integer hash -> [0,1) floats.
inline uint32_t mix32(uint32_t h) {
    h ^= h >> 16; h *= 0x7feb352du; h ^= h >> 15; h *= 0x846ca68bu; h ^= h >> 16;
    return h;
}
void jitter(int x, int y, int s, int spp, uint32_t seed, float* jx, float* jy) {
    if (spp <= 1) { *jx = 0.0f; *jy = 0.0f; return; }
    uint32_t h = seed;
    h = mix32(h ^ ((uint32_t)x * 0x9E3779B1u));
    h = mix32(h ^ ((uint32_t)y * 0x85EBCA77u));
    h = mix32(h ^ ((uint32_t)s * 0xC2B2AE3Du));
    uint32_t h2 = mix32(h + 0x68E31DA4u);
    float ux = (float)(h >> 8) * (1.0f / 16777216.0f);
    float uy = (float)(h2 >> 8) * (1.0f / 16777216.0f);
    int n = (int)floorf(sqrtf((float)spp));
    while (n * n > spp) n--;
    while ((n + 1) * (n + 1) <= spp) n++;
    if (n * n == spp) {
        int cx = s % n, cy = s / n;
        *jx = ((float)cx + ux) / (float)n;
        *jy = ((float)cy + uy) / (float)n;
    } else {
        *jx = ux; *jy = uy;
    }
}

struct CameraMats { mat4 inverse_proj_mat; };
inline CameraMats camera_mats(const rt_camera& c) {
    // glsl:366-367, 383
    mat4 proj_mat = calc_projection_matrix(c);
    mat4 view_mat = calc_view_matrix(c);
    CameraMats m;
    m.inverse_proj_mat = inverse(proj_mat * view_mat);
    return m;
}

// glsl:377-392 with the jitter extension (jx = jy = 0 reproduces the shader)
inline Ray primary_ray(const rt_camera& c, const mat4& inverse_proj_mat, int px, int py,
                       int width, int height, float jx, float jy) {
    float view_ray_x = ((float)(px - width / 2) + jx) / (float)(width / 2);
    float view_ray_y = ((float)(py - height / 2) + jy) / (float)(height / 2);
    vec4 view_ray_start(view_ray_x, view_ray_y, 0.5f, 1.0f);
    vec4 view_ray_end(view_ray_x, view_ray_y, 1.0f, 1.0f);
    vec4 world_ray_start = inverse_proj_mat * view_ray_start;
    world_ray_start = world_ray_start / world_ray_start.w;
    vec4 world_ray_end = inverse_proj_mat * view_ray_end;
    world_ray_end = world_ray_end / world_ray_end.w;
    Ray world_ray;
    world_ray.start = v3(c.position);
    world_ray.dir = normalize((world_ray_end - world_ray_start).xyz());
    return world_ray;
}

void add_stats(oracle_stats* a, const oracle_stats& b) {
    a->rays += b.rays; a->nodes += b.nodes; a->hit_nodes += b.hit_nodes; a->shadow_rays += b.shadow_rays;
    a->box_tests += b.box_tests; a->sphere_tests += b.sphere_tests; a->red_pixels += b.red_pixels;
    a->loop_iterations += b.loop_iterations;
}

}  // namespace

extern "C" {

void oracle_options_default(oracle_options* o) {
    std::memset(o, 0, sizeof(*o));
    o->spp = 1;
    o->seed = 0x5EED0001u;
}

int oracle_num_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

int oracle_render(const rt_scene* scene, const rt_camera* camera, uint8_t* out_rgba,
                  int width, int height, const oracle_options* opts) {
    oracle_options od;
    if (!opts) { oracle_options_default(&od); opts = &od; }
    if (!scene || !camera || !out_rgba || width <= 0 || height <= 0) return RT_ERR_BAD_ARG;
    if (scene->objects_count < 0 || scene->lights_count < 0 || scene->max_depth < 0) return RT_ERR_BAD_ARG;
    const int spp = opts->spp > 0 ? opts->spp : 1;
    const int y0 = opts->y0 > 0 ? opts->y0 : 0;
    const int y1 = (opts->y1 > 0 && opts->y1 <= height) ? opts->y1 : height;
    const bool as_written = opts->as_written != 0;

    std::vector<BoxXform> cache;
    if (!as_written) {
        cache.resize(scene->objects_count);
        for (int i = 0; i < scene->objects_count; i++) cache[i] = box_xform(scene->objects[i]);
    }
    CameraMats frame_mats = camera_mats(*camera);

    const bool listed = opts->pixel_list != nullptr;
    const long n_items = listed ? (long)opts->n_pixels : (long)(y1 - y0) * width;
    oracle_stats total;
    std::memset(&total, 0, sizeof(total));

#ifdef _OPENMP
    int nthreads = opts->threads > 0 ? opts->threads : omp_get_max_threads();
#pragma omp parallel num_threads(nthreads)
#endif
    {
        Invocation* inv = new Invocation;
        inv->objects = scene->objects; inv->objects_count = scene->objects_count;
        inv->lights = scene->lights; inv->lights_count = scene->lights_count;
        inv->cache = as_written ? nullptr : cache.data();
        std::memset(&inv->st, 0, sizeof(inv->st));
#ifdef _OPENMP
#pragma omp for schedule(dynamic, 64)
#endif
        for (long item = 0; item < n_items; item++) {
            int px, py;
            long out_index;
            if (listed) {
                px = opts->pixel_list[2 * item]; py = opts->pixel_list[2 * item + 1];
                out_index = item;
            } else {
                px = (int)(item % width); py = y0 + (int)(item / width);
                out_index = (long)py * width + px;
            }
            vec3 sum(0.0f);
            const uint64_t nodes_before = inv->st.nodes;
            for (int s = 0; s < spp; s++) {
                // as written the shader rebuilds the camera matrices in every invocation
                CameraMats m = as_written ? camera_mats(*camera) : frame_mats;
                float jx, jy;
                jitter(px, py, s, spp, opts->seed, &jx, &jy);
                Ray world_ray = primary_ray(*camera, m.inverse_proj_mat, px, py, width, height, jx, jy);
                vec3 col = inv->recursive_raytrace(world_ray, scene->max_depth);  // glsl:401
                sum = sum + col;
            }
            vec3 final_color = sum / (float)spp;
            // glsl:404 imageStore(vec4(final_color, 0.0)) into an RGBA8 image (main.cpp:223)
            uint8_t* o = out_rgba + 4 * out_index;
            o[0] = unorm8(final_color.x); o[1] = unorm8(final_color.y); o[2] = unorm8(final_color.z);
            o[3] = unorm8(0.0f);
            if (opts->out_nodes) opts->out_nodes[out_index] = (int32_t)(inv->st.nodes - nodes_before);
            if (opts->out_float) {
                float* f = opts->out_float + 3 * out_index;
                f[0] = final_color.x; f[1] = final_color.y; f[2] = final_color.z;
            }
        }
#ifdef _OPENMP
#pragma omp critical
#endif
        add_stats(&total, inv->st);
        delete inv;
    }
    if (opts->stats) *opts->stats = total;
    return RT_OK;
}

// glsl:199-224 lights, glsl:74-157 materials, glsl:261-321 objects
static rt_material make_material(float amb, const float* dif, float spec, float shin,
                                 float refl, float transp, float ior) {
    rt_material m;
    std::memset(&m, 0, sizeof(m));
    for (int i = 0; i < 4; i++) { m.ambient[i] = amb; m.diffuse[i] = dif[i]; m.specular[i] = spec; m.emissive[i] = 0.0f; }
    m.shininess = shin; m.reflectivity = refl; m.transparency = transp; m.refraction_index = ior;
    return m;
}

void oracle_builtin_scene(float time, rt_object* objects, rt_light* lights, rt_scene* scene) {
    const float time_scale = 0.4f;                  // glsl:236
    const float scaled_time = time * time_scale;    // glsl:237
    const float red_d[4] = {1.0f, 0.0f, 0.0f, 1.0f}, green_d[4] = {0.0f, 1.0f, 0.0f, 1.0f};
    const float blue_d[4] = {0.0f, 0.0f, 1.0f, 1.0f}, mirror_d[4] = {0.6f, 0.6f, 0.6f, 1.0f};
    const float wall_d[4] = {0.4f, 0.4f, 0.4f, 0.4f};
    rt_material red_glass = make_material(1.0f, red_d, 1.0f, 10.0f, 0.8f, 0.4f, 1.5f);      // glsl:98-108
    rt_material green_glass = make_material(1.0f, green_d, 1.0f, 10.0f, 0.4f, 0.6f, 1.5f);  // glsl:110-120
    rt_material blue_glass = make_material(1.0f, blue_d, 1.0f, 10.0f, 0.4f, 0.6f, 1.5f);    // glsl:122-132
    rt_material mirror = make_material(1.0f, mirror_d, 1.0f, 4.0f, 1.0f, 0.0f, 1.0f);       // glsl:135-145
    rt_material wall = make_material(0.5f, wall_d, 0.3f, 3.0f, 0.3f, 0.0f, 1.0f);           // glsl:147-157

    std::memset(objects, 0, 5 * sizeof(rt_object));
    auto set3 = [](float* d, float a, float b, float c) { d[0] = a; d[1] = b; d[2] = c; };
    // 0: enclosing cube (glsl:264-273)
    set3(objects[0].box_mins, -11.0f, -11.0f, -11.0f); set3(objects[0].box_maxs, 11.0f, 11.0f, 11.0f);
    objects[0].sphere_radius = -1.0f; objects[0].material = wall;
    // 1: mirror cube (glsl:275-285)
    float k = 0.5f * sinf(scaled_time * 0.5f) + 1.5f;
    set3(objects[1].box_mins, -1.0f * k, -1.0f * k, -1.0f * k); set3(objects[1].box_maxs, 1.0f * k, 1.0f * k, 1.0f * k);
    objects[1].sphere_radius = -1.0f;
    set3(objects[1].position, 0.0f, 0.0f, sinf(scaled_time * 3.0f));
    set3(objects[1].angles, 0.0f, scaled_time * 90.0f, 0.0f);
    objects[1].material = mirror;
    // 2: green slab (glsl:287-297)
    set3(objects[2].box_mins, -10.0f, -10.0f, -1.0f); set3(objects[2].box_maxs, 10.0f, 10.0f, 1.0f);
    objects[2].sphere_radius = -1.0f;
    set3(objects[2].position, 0.0f, 0.0f, -3.0f);
    set3(objects[2].angles, sinf(scaled_time * 5.0f) * 10.0f, 45.0f, 0.0f);
    objects[2].material = green_glass;
    // 3: tumbling blue box (glsl:299-309)
    set3(objects[3].box_mins, -1.0f, -1.0f, -2.0f); set3(objects[3].box_maxs, 1.0f, 1.0f, 2.0f);
    objects[3].sphere_radius = -1.0f;
    set3(objects[3].position, 3.0f, 4.0f, 1.0f);
    set3(objects[3].angles, 45.0f + scaled_time * 45.0f, 0.0f, 45.0f + scaled_time * 180.0f);
    objects[3].material = blue_glass;
    // 4: red sphere (glsl:311-319)
    objects[4].sphere_radius = 2.0f;
    set3(objects[4].position, -3.0f, 4.0f, 1.0f);
    objects[4].material = red_glass;

    std::memset(lights, 0, 3 * sizeof(rt_light));
    set3(lights[0].position, 0.1f, 0.1f, 0.1f);   // glsl:202-207
    for (int i = 0; i < 4; i++) { lights[0].ambient[i] = 0.3f; lights[0].diffuse[i] = 0.0f; lights[0].specular[i] = 0.0f; }
    set3(lights[1].position, 7.0f, 7.0f, 2.0f);   // glsl:209-214
    for (int i = 0; i < 4; i++) { lights[1].ambient[i] = 0.05f; lights[1].diffuse[i] = 1.0f; lights[1].specular[i] = 1.0f; }
    set3(lights[2].position, 3.0f, -3.0f, 4.0f);  // glsl:216-221
    for (int i = 0; i < 4; i++) { lights[2].ambient[i] = 0.05f; lights[2].diffuse[i] = red_d[i]; lights[2].specular[i] = red_d[i]; }

    scene->objects = objects; scene->objects_count = 5;   // glsl:321
    scene->lights = lights; scene->lights_count = 3;      // glsl:224
    scene->max_depth = 0;                                 // glsl:22
}

// glsl:343-364
void oracle_orbit_camera(float speed, rt_camera* c) {
    float radius = 10.0f;
    c->position[0] = radius * cosf(speed); c->position[1] = radius * sinf(speed); c->position[2] = 0.0f;
    float pitch = 0.0f, roll = 0.0f;
    float yaw = mod(1.0f * speed * (180.0f / 3.1416f), 360.0f) + 90.0f;
    c->angles[0] = pitch; c->angles[1] = yaw; c->angles[2] = roll;
    c->near_plane = 0.1f; c->far_plane = 1000.0f; c->aspect = 16.0f / 9.0f; c->v_fov = 90.0f;
}
void oracle_builtin_camera(float time, rt_camera* c) {
    float speed = time * 0.4f + 0.5f;   // glsl:343 (time_scale = 0.4, glsl:236)
    oracle_orbit_camera(speed, c);
}

void oracle_probe_ray(const rt_scene* scene, const float* start3, const float* dir3,
                      int* obj, float* t, float* p3, float* n3, int* inside) {
    Invocation* inv = new Invocation;
    inv->objects = scene->objects; inv->objects_count = scene->objects_count;
    inv->lights = scene->lights; inv->lights_count = scene->lights_count;
    inv->cache = nullptr;
    std::memset(&inv->st, 0, sizeof(inv->st));
    Ray r; r.start = v3(start3); r.dir = v3(dir3);
    Collision c = inv->get_closest_collision(r);
    if (obj) *obj = c.object_index;
    if (t) *t = c.t;
    if (p3) { p3[0] = c.p.x; p3[1] = c.p.y; p3[2] = c.p.z; }
    if (n3) { n3[0] = c.n.x; n3[1] = c.n.y; n3[2] = c.n.z; }
    if (inside) *inside = c.inside ? 1 : 0;
    delete inv;
}

void oracle_probe_pixel(const rt_scene* scene, const rt_camera* camera, int width, int height,
                        int px, int py, float* dir3, int* obj, float* t, float* p3, float* n3, int* inside) {
    CameraMats m = camera_mats(*camera);
    Ray r = primary_ray(*camera, m.inverse_proj_mat, px, py, width, height, 0.0f, 0.0f);
    if (dir3) { dir3[0] = r.dir.x; dir3[1] = r.dir.y; dir3[2] = r.dir.z; }
    float s[3] = {r.start.x, r.start.y, r.start.z}, d[3] = {r.dir.x, r.dir.y, r.dir.z};
    oracle_probe_ray(scene, s, d, obj, t, p3, n3, inside);
}

static mat4 load16(const float* m) { mat4 r; std::memcpy(r.c, m, 64); return r; }
static void store16(const mat4& m, float* o) { std::memcpy(o, m.c, 64); }
void oracle_mat4_inverse(const float* m16, float* out16) { store16(inverse(load16(m16)), out16); }
void oracle_mat4_mul(const float* a16, const float* b16, float* out16) { store16(load16(a16) * load16(b16), out16); }
void oracle_transform_matrix(const float* pos3, const float* angles3, float* out16) {
    store16(calc_transform_matrix(v3(pos3), v3(angles3)), out16);
}
void oracle_projection_matrix(const rt_camera* c, float* out16) { store16(calc_projection_matrix(*c), out16); }
void oracle_view_matrix(const rt_camera* c, float* out16) { store16(calc_view_matrix(*c), out16); }
void oracle_reflect(const float* i3, const float* n3, float* o) {
    vec3 r = reflect(v3(i3), v3(n3)); o[0] = r.x; o[1] = r.y; o[2] = r.z;
}
void oracle_refract(const float* i3, const float* n3, float eta, float* o) {
    vec3 r = refract(v3(i3), v3(n3), eta); o[0] = r.x; o[1] = r.y; o[2] = r.z;
}
uint8_t oracle_unorm8(float f) { return unorm8(f); }
void oracle_jitter(int x, int y, int s, int spp, uint32_t seed, float* jx, float* jy) { jitter(x, y, s, spp, seed, jx, jy); }

}  // extern "C"
