/*
 * rt_scene_io.cpp — scene / camera description on the wire.
 *
 * The reference hard-codes its scene in shader globals (glsl:74-157 materials,
 * glsl:199-224 lights, glsl:261-321 objects, glsl:334-364 camera, glsl:22 depth).
 * This is the same data as a line-oriented text file, one record per line, the
 * fields in the order of the shader structs (include/rt_b200.h):
 *
 *   rtscene 1
 *   depth  D                                    MAX_RAYTRACE_DEPTH (glsl:22)
 *   camera px py pz  pitch yaw roll  v_fov aspect near far         (glsl:36-50)
 *   light  px py pz  amb[4]  dif[4]  spec[4]                       (glsl:190-196)
 *   object mins[3] maxs[3] radius  pos[3] angles[3]  ambient[4] diffuse[4] specular[4]
 *          shininess emissive[4] reflectivity transparency refraction_index   (glsl:244-258, 56-69)
 *   # comment
 *
 * Floats are written with 9 significant digits, so save -> load reproduces every
 * FP32 bit pattern.  Null conventions are the shader's (glsl:178-179): an
 * all-zero box is "no box", radius -1 is "no sphere".
 */
#include <cerrno>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>
#include <string>
#include <vector>

#include "../../include/rt_b200.h"

extern "C" int rt_set_error_(int code, const char* msg);   // rt_host.cpp

struct rt_scene_file {
    std::vector<rt_object> objects;
    std::vector<rt_light> lights;
    rt_scene scene;
    rt_camera camera;
    int has_camera;
};

namespace {

int err(int code, const std::string& m) { return rt_set_error_(code, m.c_str()); }

// Parses exactly n floats from p; returns false on a malformed or missing number.
bool floats(const char*& p, float* dst, int n) {
    for (int i = 0; i < n; i++) {
        char* end = nullptr;
        errno = 0;
        const float v = strtof(p, &end);
        if (end == p || errno == ERANGE) return false;
        dst[i] = v;
        p = end;
    }
    return true;
}

bool only_space_left(const char* p) {
    while (*p == ' ' || *p == '\t' || *p == '\r' || *p == '\n') p++;
    return *p == '\0' || *p == '#';
}

void put(FILE* f, const float* v, int n) {
    for (int i = 0; i < n; i++) fprintf(f, " %.9g", (double)v[i]);
}

void fix_pointers(rt_scene_file* sf) {
    sf->scene.objects = sf->objects.empty() ? nullptr : sf->objects.data();
    sf->scene.objects_count = (int32_t)sf->objects.size();
    sf->scene.lights = sf->lights.empty() ? nullptr : sf->lights.data();
    sf->scene.lights_count = (int32_t)sf->lights.size();
}

}  // namespace

extern "C" {

int rt_scene_save(const char* path, const rt_scene* scene, const rt_camera* camera) {
    if (!path || !scene) return err(RT_ERR_BAD_ARG, "rt_scene_save: path or scene is NULL");
    if (scene->objects_count < 0 || scene->lights_count < 0 || (scene->objects_count > 0 && !scene->objects) ||
        (scene->lights_count > 0 && !scene->lights))
        return err(RT_ERR_BAD_ARG, "rt_scene_save: bad objects / lights");
    FILE* f = fopen(path, "w");
    if (!f) return err(RT_ERR_BAD_ARG, std::string("rt_scene_save: cannot open ") + path);
    fprintf(f, "rtscene 1\n# fields in the order of the shader structs (include/rt_b200.h)\ndepth %d\n", scene->max_depth);
    if (camera) {
        fprintf(f, "camera");
        put(f, camera->position, 3); put(f, camera->angles, 3);
        put(f, &camera->v_fov, 1); put(f, &camera->aspect, 1); put(f, &camera->near_plane, 1); put(f, &camera->far_plane, 1);
        fprintf(f, "\n");
    }
    for (int j = 0; j < scene->lights_count; j++) {
        fprintf(f, "light");
        put(f, reinterpret_cast<const float*>(&scene->lights[j]), 15);
        fprintf(f, "\n");
    }
    for (int i = 0; i < scene->objects_count; i++) {
        fprintf(f, "object");
        put(f, reinterpret_cast<const float*>(&scene->objects[i]), 33);
        fprintf(f, "\n");
    }
    const bool bad = ferror(f) != 0;
    if (fclose(f) != 0 || bad) return err(RT_ERR_BAD_ARG, std::string("rt_scene_save: write failed: ") + path);
    return RT_OK;
}

int rt_scene_load(const char* path, rt_scene_file** out) {
    if (!path || !out) return err(RT_ERR_BAD_ARG, "rt_scene_load: path or out is NULL");
    *out = nullptr;
    FILE* f = fopen(path, "r");
    if (!f) return err(RT_ERR_BAD_ARG, std::string("rt_scene_load: cannot open ") + path);
    rt_scene_file* sf = new (std::nothrow) rt_scene_file;
    if (!sf) { fclose(f); return err(RT_ERR_NOMEM, "rt_scene_load: out of memory"); }
    memset(&sf->scene, 0, sizeof(sf->scene));
    memset(&sf->camera, 0, sizeof(sf->camera));
    sf->has_camera = 0;
    std::string line;
    char buf[4096];
    int lineno = 0;
    bool header = false;
    int rc = RT_OK;
    std::string why;
    while (rc == RT_OK) {
        // a record may be longer than buf: keep reading until the newline
        line.clear();
        bool got = false;
        while (fgets(buf, sizeof(buf), f)) {
            got = true;
            line += buf;
            if (!line.empty() && line.back() == '\n') break;
        }
        if (!got) break;
        lineno++;
        const char* p = line.c_str();
        while (*p == ' ' || *p == '\t') p++;
        if (*p == '\0' || *p == '\n' || *p == '\r' || *p == '#') continue;
        char key[16] = "";
        int used = 0;
        if (sscanf(p, "%15s%n", key, &used) != 1) continue;
        p += used;
        if (!header) {
            float ver[1];
            if (strcmp(key, "rtscene") != 0 || !floats(p, ver, 1) || ver[0] != 1.0f) {
                rc = RT_ERR_BAD_ARG; why = "expected header 'rtscene 1'";
            }
            header = true;
            continue;
        }
        if (strcmp(key, "depth") == 0) {
            float d[1];
            if (!floats(p, d, 1) || d[0] < 0.0f || d[0] != (float)(int)d[0]) { rc = RT_ERR_BAD_ARG; why = "bad depth"; }
            else sf->scene.max_depth = (int32_t)d[0];
        } else if (strcmp(key, "camera") == 0) {
            if (!floats(p, reinterpret_cast<float*>(&sf->camera), 10)) { rc = RT_ERR_BAD_ARG; why = "camera needs 10 numbers"; }
            else sf->has_camera = 1;
        } else if (strcmp(key, "light") == 0) {
            rt_light L;
            if (!floats(p, reinterpret_cast<float*>(&L), 15)) { rc = RT_ERR_BAD_ARG; why = "light needs 15 numbers"; }
            else sf->lights.push_back(L);
        } else if (strcmp(key, "object") == 0) {
            rt_object o;
            if (!floats(p, reinterpret_cast<float*>(&o), 33)) { rc = RT_ERR_BAD_ARG; why = "object needs 33 numbers"; }
            else sf->objects.push_back(o);
        } else {
            rc = RT_ERR_BAD_ARG; why = std::string("unknown record '") + key + "'";
        }
        if (rc == RT_OK && !only_space_left(p)) { rc = RT_ERR_BAD_ARG; why = "trailing characters"; }
    }
    fclose(f);
    if (rc == RT_OK && !header) { rc = RT_ERR_BAD_ARG; why = "empty file"; }
    if (rc == RT_OK && (int)sf->lights.size() > RT_MAX_LIGHTS) { rc = RT_ERR_BAD_ARG; why = "too many lights"; }
    if (rc != RT_OK) {
        delete sf;
        char where[64];
        snprintf(where, sizeof(where), ":%d: ", lineno);
        return err(rc, std::string("rt_scene_load: ") + path + where + why);
    }
    fix_pointers(sf);
    *out = sf;
    return RT_OK;
}

const rt_scene* rt_scene_file_scene(const rt_scene_file* sf) { return sf ? &sf->scene : nullptr; }
const rt_camera* rt_scene_file_camera(const rt_scene_file* sf) { return (sf && sf->has_camera) ? &sf->camera : nullptr; }
void rt_scene_file_free(rt_scene_file* sf) { delete sf; }

}  // extern "C"
