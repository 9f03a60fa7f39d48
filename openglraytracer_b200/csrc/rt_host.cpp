/*
 * rt_host.cpp — headless host side of the C ABI (include/rt_b200.h).
 *
 * Replaces the reference's GL host path for the hot path only: context +
 * texture (main.cpp:47-76, 152-159) -> rt_context with device buffers;
 * glUniform1f/glDispatchCompute/glFinish (main.cpp:220-238) -> rt_render*.
 * The host prologue hoists what the shader recomputes per pixel / per test:
 * inverse(proj*view) (glsl:366-383) once per frame, the box transforms
 * (glsl:650-652, 718) once per object.
 *
 * No CPU fallback: without a usable CUDA device every render call fails with
 * RT_ERR_NO_DEVICE / RT_ERR_CUDA.
 */
#include <cuda_runtime.h>

#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <vector>

#include "../../include/rt_b200.h"
#include "rt_device.h"
#include "rt_hostmath.h"
#include "rt_prologue_math.h"

static_assert(RT_MAX_DEPTH == RT_MAX_DEPTH_DEV, "depth limits out of sync");
static_assert(sizeof(rt_material) == 20 * 4, "rt_material must be 20 words (glsl:56-69)");
static_assert(sizeof(rt_object) == 33 * 4, "rt_object must be 33 words (glsl:244-258)");
static_assert(sizeof(rt_light) == 15 * 4, "rt_light must be 15 words (glsl:190-196)");
static_assert(sizeof(rt_camera) == 10 * 4, "rt_camera must be 10 words (glsl:36-50)");
static_assert(sizeof(RtCold) == 64 * 4, "cold record must be 256 bytes");
static_assert(RT_NOALLOC_MAX_OBJECTS == RT_NOALLOC_MAX_OBJECTS_DEV && RT_NOALLOC_MAX_FRAMES == RT_NOALLOC_MAX_FRAMES_DEV &&
              RT_NOALLOC_SCRATCH_BYTES == RT_NOALLOC_SCRATCH_BYTES_DEV, "no-alloc limits out of sync");

namespace rt_fast {
cudaError_t launch_block_noalloc(const RtParams& P, const RtCameraDev* cams_host, int device, int sm_count,
                                 cudaStream_t stream, const char** why);
cudaError_t launch(const RtParams& P, int sm_count, bool count, cudaStream_t stream);
cudaError_t configure(int device, int* max_smem_optin);
size_t resident_limit_bytes(int max_smem_optin, bool cull);
}
namespace rt_prologue {
int max_grid(int sm_count);
size_t scratch_bytes(int max_ctas);
cudaError_t run(const rt_object* d_objs, int n_obj, int32_t* d_box_obj, int32_t* d_sph_obj, RtCold* d_cold,
                float* d_box_fast, float* d_box_cull, float* d_sph, const RtLightDev* d_lights, int n_lit, const int* lit,
                float light_reach, float* d_obj_lrec, void* d_scratch, int sm_count, cudaStream_t st);
}
namespace rt_groups {
size_t scratch_bytes(int n_hot, int sm_count);
cudaError_t run(const float* d_box_cull, const float* d_sph, int n_box, int n_sph, const float* d_extent, float* d_grp, float* d_srec,
                uint32_t* d_sidx, void* d_scratch, int sm_count, cudaStream_t st);
}
namespace rt_strict {
cudaError_t launch(const RtParams& P, int sm_count, bool count, cudaStream_t stream);
cudaError_t configure(int device, int* max_smem_optin);
size_t resident_limit_bytes(int max_smem_optin, bool cull);
}

namespace {

thread_local char g_err[512] = "";
std::atomic<uint64_t> g_launches{0};

int fail(int code, const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
    return code;
}

#define CU(call)                                                                          \
    do {                                                                                  \
        cudaError_t e_ = (call);                                                          \
        if (e_ != cudaSuccess)                                                            \
            return fail(e_ == cudaErrorNoDevice || e_ == cudaErrorInsufficientDriver ||   \
                                e_ == cudaErrorNoKernelImageForDevice                     \
                            ? RT_ERR_NO_DEVICE                                            \
                            : RT_ERR_CUDA,                                                \
                        "%s failed: %s", #call, cudaGetErrorString(e_));                  \
    } while (0)

template <typename T>
struct DevBuf {
    T* p = nullptr;
    size_t cap = 0;   // elements
    cudaError_t reserve(size_t n) {
        if (n <= cap) return cudaSuccess;
        if (p) cudaFree(p);
        p = nullptr; cap = 0;
        size_t want = n + n / 4 + 16;
        cudaError_t e = cudaMalloc((void**)&p, want * sizeof(T));
        if (e == cudaSuccess) cap = want;
        return e;
    }
    void release() { if (p) cudaFree(p); p = nullptr; cap = 0; }
};

struct HostScene {
    std::vector<float> box_fast, box_strict;   // float4 records
    std::vector<float> box_cull;               // (world-space centre, padded squared circumradius) per box
    std::vector<float> obj_lrec;               // per 64-object tile x lit light: (L - c, |L - c|^2 - R'^2)
    int n_lit = 0;                             // lit lights with records (0: more than 3, or none)
    std::vector<float> sph;                    // (c, r*r)
    std::vector<int32_t> box_obj, sph_obj;
    std::vector<int32_t> kind, hot_index;      // per original object
    int lit[3] = {0, 0, 0};                    // indices of the lit lights that have records
    double light_reach = 1.0;                  // max |light position| (part of the scene extent)
    float extent_f = 1.0f;                     // scene extent E, rounded up
    int n_lit_total = 0;                       // lit lights, any number
    float amb_sum[4] = {0, 0, 0, 0};           // sum of the lights' ambient terms in light order (glsl:801)
    std::vector<RtCold> cold;
    std::vector<RtLightDev> lights;
    int max_depth = 0;
};

}  // namespace

struct rt_context {
    int device = 0;
    int sm_count = 0;
    int max_smem_optin = 0;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev[4] = {nullptr, nullptr, nullptr, nullptr};
    // Ordering between the context's own stream (scene uploads, prologue kernels) and a caller's stream
    // (rt_options.stream): a render on another stream waits for ev_upload; the next upload waits (on the host,
    // it rewrites the page-locked staging block and the device arena) for ev_render of the last such render.
    cudaEvent_t ev_upload = nullptr, ev_render = nullptr;
    bool upload_recorded = false, render_recorded = false;
    DevBuf<float> d_box_fast, d_box_strict, d_box_cull, d_obj_lrec, d_sph;
    DevBuf<int32_t> d_box_obj, d_sph_obj;
    DevBuf<RtCold> d_cold;
    DevBuf<RtLightDev> d_lights;
    DevBuf<RtCameraDev> d_cams;
    DevBuf<uint8_t> d_out;
    DevBuf<RtCounters> d_cnt;
    DevBuf<unsigned long long> d_tile;
    DevBuf<rt_object> d_objs;                 // device prologue: the raw objects
    DevBuf<uint8_t> d_scratch;                // device prologue: counts, extent, barrier words
    size_t n_objects_dev = 0;
    DevBuf<float> d_grp, d_srec;              // groups of neighbouring objects (rt_groups.cu): spheres, sorted bounding records
    DevBuf<uint32_t> d_sidx;                  // hot-list positions of the sorted records
    DevBuf<uint8_t> d_gscratch;
    int n_groups = 0;                         // 0: none built for the uploaded scene
    DevBuf<float> d_brec;                     // ... and the bounding records of each list's first RT_BLOCK_REC_CAP entries
    DevBuf<uint32_t> d_blist;                 // block kernel: shadow-candidate lists, one slice per resident CTA
    DevBuf<uint32_t> d_blklog;                // block kernel, RT_BLOCK_LOG=path: per-block timing records of the instrumented twin
    int prologue_mode = RT_PROLOGUE_AUTO;
    bool strict_ready = false;                // d_box_strict holds the current scene
    std::vector<rt_object> objs_copy;         // strict render after a device prologue: fetched back from d_objs
    std::vector<rt_light> lights_copy;
    int depth_copy = 0;
    uint8_t* h_pinned = nullptr;
    size_t h_pinned_cap = 0;
    // Host prologue: every record array is packed into one page-locked staging block and goes up in ONE copy into
    // one device arena (ten small pageable copies and a second stream synchronisation were 4 % of the built-in
    // frame's end-to-end time).  The render reads the arrays through these pointers, whichever prologue wrote them.
    DevBuf<uint8_t> d_arena;
    uint8_t* h_stage = nullptr;
    size_t h_stage_cap = 0;
    const float *p_box_fast = nullptr, *p_box_strict = nullptr, *p_box_cull = nullptr, *p_obj_lrec = nullptr, *p_sph = nullptr;
    const int32_t *p_box_obj = nullptr, *p_sph_obj = nullptr;
    const RtCold* p_cold = nullptr;
    const RtLightDev* p_lights = nullptr;
    const unsigned int* p_extent = nullptr;
    int room_skip = 0;                        // hot box 0 holds every lit light (host prologue only)
    // uploaded scene
    bool have_scene = false;
    int n_box = 0, n_sph = 0, n_lights = 0, max_depth = 0, n_lit = 0;
    int n_lit_total = 0, lit[3] = {0, 0, 0};
    float amb_sum[4] = {0, 0, 0, 0};
    float4 lit_pos[3], lit_dif[3], lit_spc[3];
};

namespace {

// The lights: device records, the lit ones (a light with zero diffuse and specular adds 0 shadowed or not, glsl:830-832),
// the ambient sum (glsl:801), max |light position| (part of the scene extent).
void classify_lights(const rt_scene* scene, HostScene* hs) {
    hs->max_depth = scene->max_depth;
    hs->lights.resize(scene->lights_count > 0 ? scene->lights_count : 0);
    std::vector<int> lit;
    for (int j = 0; j < scene->lights_count; j++) {
        const rt_light& L = scene->lights[j];
        RtLightDev& d = hs->lights[j];
        memset(&d, 0, sizeof(d));
        memcpy(d.position, L.position, 12);
        memcpy(d.ambient, L.ambient, 16);
        memcpy(d.diffuse, L.diffuse, 16);
        memcpy(d.specular, L.specular, 16);
        bool is_lit = false;
        for (int k = 0; k < 4; k++) is_lit = is_lit || L.diffuse[k] != 0.0f || L.specular[k] != 0.0f;
        d.position[3] = is_lit ? 1.0f : 0.0f;
        if (is_lit) lit.push_back(j);
    }
    // per-light records exist for 1..3 lit lights (one shadow batch); otherwise the generic path runs
    hs->n_lit = (lit.size() >= 1 && lit.size() <= 3) ? (int)lit.size() : 0;
    hs->n_lit_total = (int)lit.size();
    for (int k = 0; k < 4; k++) {
        float acc = 0.0f;
        for (const RtLightDev& L : hs->lights) acc = acc + L.ambient[k];
        hs->amb_sum[k] = acc;
    }
    for (int s2 = 0; s2 < 3; s2++) hs->lit[s2] = s2 < hs->n_lit ? lit[s2] : 0;
    hs->light_reach = 1.0;
    for (const RtLightDev& L : hs->lights) {
        const float zero[3] = {L.position[0], L.position[1], L.position[2]};
        hs->light_reach = std::max(hs->light_reach, rtpm::reach(zero, 0.0));
    }
}

// Classification (glsl:749, 759) and the hot lists in original order (host prologue).
void classify_scene(const rt_scene* scene, HostScene* hs) {
    const int n = scene->objects_count;
    hs->kind.assign(n > 0 ? n : 0, RT_KIND_SKIP);
    hs->hot_index.assign(n > 0 ? n : 0, -1);
    hs->box_obj.clear(); hs->sph_obj.clear();
    for (int i = 0; i < n; i++) {
        const int k = rtpm::object_kind(scene->objects[i]);
        hs->kind[i] = k;
        if (k == RT_KIND_BOX) { hs->hot_index[i] = (int32_t)hs->box_obj.size(); hs->box_obj.push_back(i); }
        else if (k == RT_KIND_SPHERE) { hs->hot_index[i] = (int32_t)hs->sph_obj.size(); hs->sph_obj.push_back(i); }
    }
    classify_lights(scene, hs);
}

// Host prologue: per-object matrices and records (rt_prologue_math.h), then the padded bounding records.
int build_host_scene(const rt_scene* scene, HostScene* hs) {
    classify_scene(scene, hs);
    const int n = scene->objects_count;
    const size_t nb = hs->box_obj.size(), ns = hs->sph_obj.size();
    hs->cold.resize(n > 0 ? n : 0);
    hs->box_fast.assign(nb * 12, 0.0f);
    hs->box_strict.assign(nb * 20, 0.0f);
    hs->box_cull.assign(nb * 4, 0.0f);
    hs->sph.assign(ns * 4, 0.0f);
    for (int i = 0; i < n; i++) {
        const int k = hs->kind[i], h = hs->hot_index[i];
        float dummy[20];
        rtpm::object_records(scene->objects[i], k, h, &hs->cold[i],
                             k == RT_KIND_BOX ? &hs->box_fast[(size_t)h * 12] : dummy,
                             k == RT_KIND_BOX ? &hs->box_strict[(size_t)h * 20] : dummy,
                             k == RT_KIND_BOX ? &hs->box_cull[(size_t)h * 4] : dummy,
                             k == RT_KIND_SPHERE ? &hs->sph[(size_t)h * 4] : dummy);
    }
    double extent = hs->light_reach;
    for (size_t b = 0; b < nb; b++) extent = std::max(extent, rtpm::reach(&hs->box_cull[4 * b], std::fabs((double)hs->box_cull[4 * b + 3])));
    for (size_t j = 0; j < ns; j++) extent = std::max(extent, rtpm::reach(&hs->sph[4 * j], std::sqrt(std::fabs((double)hs->sph[4 * j + 3]))));
    if (hs->n_lit > 0) {
        const int n_hot = (int)(nb + ns);
        const int tiles = (n_hot + 63) / 64;
        hs->obj_lrec.assign((size_t)tiles * hs->n_lit * 64 * 4, 0.0f);
        for (size_t k = 3; k < hs->obj_lrec.size(); k += 4) hs->obj_lrec[k] = 1e30f;   // padding: never passes
        for (int h = 0; h < n_hot; h++) {
            const bool is_b = h < (int)nb;
            const float* c3 = is_b ? &hs->box_cull[4 * h] : &hs->sph[4 * (h - (int)nb)];
            const double rad = is_b ? (double)c3[3] : std::sqrt((double)c3[3]);
            for (int s2 = 0; s2 < hs->n_lit; s2++)
                rtpm::light_record(c3, rad, is_b, hs->lights[hs->lit[s2]].position, extent,
                                   &hs->obj_lrec[(((size_t)(h / 64) * hs->n_lit + s2) * 64 + (h % 64)) * 4]);
        }
    }
    for (size_t b = 0; b < nb; b++) hs->box_cull[4 * b + 3] = rtpm::padded_radius2(hs->box_cull[4 * b + 3], extent);
    hs->extent_f = (float)(extent * 1.000001);      // the kernel's packet tests scale their pads with it
    return RT_OK;
}

int validate_scene(const rt_scene* scene) {
    if (!scene) return fail(RT_ERR_BAD_ARG, "scene is NULL");
    if (scene->objects_count < 0 || (scene->objects_count > 0 && !scene->objects))
        return fail(RT_ERR_BAD_ARG, "bad objects / objects_count");
    if (scene->lights_count < 0 || scene->lights_count > RT_MAX_LIGHTS || (scene->lights_count > 0 && !scene->lights))
        return fail(RT_ERR_BAD_ARG, "bad lights / lights_count (max %d)", RT_MAX_LIGHTS);
    if (scene->max_depth < 0 || scene->max_depth > RT_MAX_DEPTH)
        return fail(RT_ERR_BAD_ARG, "max_depth %d outside 0..%d", scene->max_depth, RT_MAX_DEPTH);
    return RT_OK;
}

}  // namespace

// =========================================================================
// C ABI
// =========================================================================
extern "C" {

int rt_abi_version(void) { return RT_ABI_VERSION; }
/* shared error sink for the other translation units of the library (rt_scene_io.cpp); not in the header */
int rt_set_error_(int code, const char* msg) { return fail(code, "%s", msg ? msg : ""); }
const char* rt_last_error_string(void) { return g_err; }
uint64_t rt_kernel_launch_count(void) { return g_launches.load(); }

void rt_options_default(rt_options* o) {
    if (!o) return;
    memset(o, 0, sizeof(*o));
    o->spp = 1;
    o->seed = 0x5EED0001u;
    o->device = -1;
}

int rt_create(rt_context** out_ctx, int device) {
    if (!out_ctx) return fail(RT_ERR_BAD_ARG, "out_ctx is NULL");
    *out_ctx = nullptr;
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0)
        return fail(RT_ERR_NO_DEVICE, "no CUDA device (%s); this library has no CPU fallback",
                    e != cudaSuccess ? cudaGetErrorString(e) : "device count 0");
    if (device < 0) CU(cudaGetDevice(&device));
    if (device >= ndev) return fail(RT_ERR_BAD_ARG, "device %d out of range (%d devices)", device, ndev);
    CU(cudaSetDevice(device));
    rt_context* c = new (std::nothrow) rt_context;
    if (!c) return fail(RT_ERR_NOMEM, "out of host memory");
    c->device = device;
    CU(cudaDeviceGetAttribute(&c->sm_count, cudaDevAttrMultiProcessorCount, device));
    CU(rt_fast::configure(device, &c->max_smem_optin));
    CU(rt_strict::configure(device, &c->max_smem_optin));
    CU(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
    for (int i = 0; i < 4; i++) CU(cudaEventCreate(&c->ev[i]));
    CU(cudaEventCreateWithFlags(&c->ev_upload, cudaEventDisableTiming));
    CU(cudaEventCreateWithFlags(&c->ev_render, cudaEventDisableTiming));
    CU(c->d_cnt.reserve(1));
    CU(c->d_tile.reserve(2));
    *out_ctx = c;
    return RT_OK;
}

int rt_destroy(rt_context* c) {
    if (!c) return RT_OK;
    cudaSetDevice(c->device);
    if (c->stream) cudaStreamSynchronize(c->stream);
    c->d_box_fast.release(); c->d_box_strict.release(); c->d_box_cull.release(); c->d_obj_lrec.release(); c->d_sph.release();
    c->d_box_obj.release(); c->d_sph_obj.release(); c->d_cold.release();
    c->d_objs.release(); c->d_scratch.release(); c->d_blist.release(); c->d_brec.release(); c->d_blklog.release();
    c->d_grp.release(); c->d_srec.release(); c->d_sidx.release(); c->d_gscratch.release();
    c->d_lights.release(); c->d_cams.release(); c->d_out.release(); c->d_cnt.release(); c->d_tile.release();
    if (c->h_pinned) cudaFreeHost(c->h_pinned);
    if (c->h_stage) cudaFreeHost(c->h_stage);
    c->d_arena.release();
    for (int i = 0; i < 4; i++) if (c->ev[i]) cudaEventDestroy(c->ev[i]);
    if (c->render_recorded) cudaEventSynchronize(c->ev_render);
    if (c->ev_upload) cudaEventDestroy(c->ev_upload);
    if (c->ev_render) cudaEventDestroy(c->ev_render);
    if (c->stream) cudaStreamDestroy(c->stream);
    delete c;
    return RT_OK;
}

int rt_synchronize(rt_context* c) {
    if (!c) return fail(RT_ERR_BAD_ARG, "ctx is NULL");
    CU(cudaSetDevice(c->device));
    CU(cudaStreamSynchronize(c->stream));
    return RT_OK;
}

int rt_host_alloc(void** ptr, size_t bytes) {
    if (!ptr || bytes == 0) return fail(RT_ERR_BAD_ARG, "bad rt_host_alloc arguments");
    CU(cudaHostAlloc(ptr, bytes, cudaHostAllocPortable));
    return RT_OK;
}
int rt_host_free(void* ptr) {
    if (ptr) CU(cudaFreeHost(ptr));
    return RT_OK;
}

// Groups of 32 neighbouring objects with a bounding sphere each (rt_groups.cu), from the records the prologue left on
// the device; one launch on the context's stream, behind the prologue, in front of the upload event.
static int build_groups(rt_context* c) {
    static const int min_objects = getenv("RT_GROUPS_MIN") ? atoi(getenv("RT_GROUPS_MIN")) : RT_GROUPS_MIN_OBJECTS;
    const int n_hot = c->n_box + c->n_sph;
    c->n_groups = 0;
    if (n_hot < min_objects || n_hot < 1) return RT_OK;
    const size_t ng = ((size_t)n_hot + RT_GROUP_SIZE - 1) / RT_GROUP_SIZE;
    CU(c->d_grp.reserve(ng * 4));
    CU(c->d_srec.reserve(ng * RT_GROUP_SIZE * 4));
    CU(c->d_sidx.reserve(ng * RT_GROUP_SIZE));
    CU(c->d_gscratch.reserve(rt_groups::scratch_bytes(n_hot, c->sm_count)));
    CU(rt_groups::run(c->p_box_cull, c->p_sph, c->n_box, c->n_sph, reinterpret_cast<const float*>(c->p_extent), c->d_grp.p, c->d_srec.p,
                      c->d_sidx.p, c->d_gscratch.p, c->sm_count, c->stream));
    g_launches += 1;
    c->n_groups = (int)ng;
    return RT_OK;
}

// Host prologue + upload of every record (all kernels, incl. strict).
static int upload_host(rt_context* c, const rt_scene* scene) {
    HostScene hs;
    int rc = build_host_scene(scene, &hs);
    if (rc != RT_OK) return rc;
    CU(cudaSetDevice(c->device));
    // The previous frame may still be reading these buffers -- on the context's stream or on a caller's.
    CU(cudaStreamSynchronize(c->stream));
    if (c->render_recorded) CU(cudaEventSynchronize(c->ev_render));
    // layout of the arena: 256-byte aligned sections, 256 bytes of slack after each (pipelined record loads may read
    // a few records past the end of a list; what they read is never used)
    struct Sec { const void* src; size_t bytes; size_t off; };
    Sec sec[10] = {{&hs.extent_f, 4, 0}, {hs.box_fast.data(), hs.box_fast.size() * 4, 0}, {hs.box_strict.data(), hs.box_strict.size() * 4, 0},
                   {hs.box_cull.data(), hs.box_cull.size() * 4, 0}, {hs.obj_lrec.data(), hs.obj_lrec.size() * 4, 0},
                   {hs.sph.data(), hs.sph.size() * 4, 0}, {hs.box_obj.data(), hs.box_obj.size() * 4, 0},
                   {hs.sph_obj.data(), hs.sph_obj.size() * 4, 0}, {hs.cold.data(), hs.cold.size() * sizeof(RtCold), 0},
                   {hs.lights.data(), hs.lights.size() * sizeof(RtLightDev), 0}};
    size_t total = 0;
    for (Sec& x : sec) { x.off = total; total += (x.bytes + 256 + 255) / 256 * 256; }
    CU(c->d_arena.reserve(total));
    if (total > c->h_stage_cap) {
        if (c->h_stage) cudaFreeHost(c->h_stage);
        c->h_stage = nullptr; c->h_stage_cap = 0;
        CU(cudaHostAlloc((void**)&c->h_stage, total + total / 4, cudaHostAllocDefault));
        c->h_stage_cap = total + total / 4;
    }
    // the stream is idle (synchronised above), so the staging block of the previous upload has been consumed
    for (const Sec& x : sec) if (x.bytes) memcpy(c->h_stage + x.off, x.src, x.bytes);
    CU(cudaMemcpyAsync(c->d_arena.p, c->h_stage, total, cudaMemcpyHostToDevice, c->stream));
    CU(cudaEventRecord(c->ev_upload, c->stream));
    c->upload_recorded = true;
    uint8_t* a = c->d_arena.p;
    c->p_extent = reinterpret_cast<const unsigned int*>(a + sec[0].off);
    c->p_box_fast = reinterpret_cast<const float*>(a + sec[1].off);
    c->p_box_strict = reinterpret_cast<const float*>(a + sec[2].off);
    c->p_box_cull = reinterpret_cast<const float*>(a + sec[3].off);
    c->p_obj_lrec = reinterpret_cast<const float*>(a + sec[4].off);
    c->p_sph = reinterpret_cast<const float*>(a + sec[5].off);
    c->p_box_obj = reinterpret_cast<const int32_t*>(a + sec[6].off);
    c->p_sph_obj = reinterpret_cast<const int32_t*>(a + sec[7].off);
    c->p_cold = reinterpret_cast<const RtCold*>(a + sec[8].off);
    c->p_lights = reinterpret_cast<const RtLightDev*>(a + sec[9].off);
    c->n_box = (int)hs.box_obj.size();
    c->n_sph = (int)hs.sph_obj.size();
    c->n_lights = (int)hs.lights.size();
    c->n_lit = hs.n_lit;
    c->n_lit_total = hs.n_lit_total;
    memcpy(c->lit, hs.lit, sizeof(c->lit)); memcpy(c->amb_sum, hs.amb_sum, sizeof(c->amb_sum));
    for (int k = 0; k < 3; k++) {
        const bool on = k < hs.n_lit_total && hs.n_lit_total <= 3;
        const RtLightDev z = {{0, 0, 0, 0}, {0, 0, 0, 0}, {0, 0, 0, 0}, {0, 0, 0, 0}};
        const RtLightDev& L = on ? hs.lights[hs.lit[k]] : z;
        c->lit_pos[k] = make_float4(L.position[0], L.position[1], L.position[2], 0.0f);
        c->lit_dif[k] = make_float4(L.diffuse[0], L.diffuse[1], L.diffuse[2], L.diffuse[3]);
        c->lit_spc[k] = make_float4(L.specular[0], L.specular[1], L.specular[2], L.specular[3]);
    }
    c->max_depth = hs.max_depth;
    c->have_scene = true;
    c->strict_ready = true;
    {
        const int rc2 = build_groups(c);
        if (rc2 != RT_OK) return rc2;
        if (c->n_groups > 0) CU(cudaEventRecord(c->ev_upload, c->stream));       // the upload now ends with the group kernel
    }
    // does the first box hold every lit light, 1e-3 (of the half extent) away from its faces?  (all_objects: room_skip)
    c->room_skip = 0;
    if (!hs.box_obj.empty() && hs.n_lit_total > 0) {
        bool all_in = true;
        const float* r = &hs.box_fast[0];
        for (const RtLightDev& L : hs.lights) {
            if (L.position[3] == 0.0f) continue;          // unlit: casts no traversed shadow ray
            for (int a = 0; a < 3; a++) {
                const double v = (double)r[4 * a] * L.position[0] + (double)r[4 * a + 1] * L.position[1] +
                                 (double)r[4 * a + 2] * L.position[2] + (double)r[4 * a + 3];
                all_in = all_in && std::fabs(v) < 0.999;
            }
        }
        c->room_skip = all_in ? 1 : 0;
    }
    return RT_OK;
}

// Device prologue (rt_prologue.cu): the raw objects go up in one copy straight from the caller's array, one kernel
// classifies them, assigns the hot-list positions and builds every record in HBM.  Fast kernel only; a later strict
// render fetches the objects back and re-runs the host prologue.
static int upload_device(rt_context* c, const rt_scene* scene) {
    HostScene hs;
    classify_lights(scene, &hs);
    const size_t n = (size_t)(scene->objects_count > 0 ? scene->objects_count : 0);
    CU(cudaSetDevice(c->device));
    // The previous frame may still be reading these buffers -- on the context's stream or on a caller's.
    CU(cudaStreamSynchronize(c->stream));
    if (c->render_recorded) CU(cudaEventSynchronize(c->ev_render));
    const int grid_cap = rt_prologue::max_grid(c->sm_count);
    CU(c->d_objs.reserve(n + 1));
    CU(c->d_scratch.reserve(rt_prologue::scratch_bytes(grid_cap)));
    CU(c->d_box_fast.reserve(n * 12 + 4));
    CU(c->d_box_cull.reserve(n * 4 + 4));
    CU(c->d_obj_lrec.reserve(((n + 63) / 64) * (size_t)(hs.n_lit > 0 ? hs.n_lit : 1) * 64 * 4 + 4));
    CU(c->d_sph.reserve(n * 4 + 4));
    CU(c->d_box_obj.reserve(n + 1));
    CU(c->d_sph_obj.reserve(n + 1));
    CU(c->d_cold.reserve(n + 1));
    CU(c->d_lights.reserve(hs.lights.size() + 1));
    if (n) CU(cudaMemcpyAsync(c->d_objs.p, scene->objects, n * sizeof(rt_object), cudaMemcpyHostToDevice, c->stream));
    if (!hs.lights.empty())
        CU(cudaMemcpyAsync(c->d_lights.p, hs.lights.data(), hs.lights.size() * sizeof(RtLightDev), cudaMemcpyHostToDevice, c->stream));
    CU(rt_prologue::run(c->d_objs.p, (int)n, c->d_box_obj.p, c->d_sph_obj.p, c->d_cold.p, c->d_box_fast.p, c->d_box_cull.p,
                        c->d_sph.p, c->d_lights.p, hs.n_lit, hs.lit, (float)hs.light_reach, c->d_obj_lrec.p, c->d_scratch.p,
                        c->sm_count, c->stream));
    g_launches += 1;
    unsigned int back[3] = {0, 0, 0};      // n_box, n_sph, extent bits
    CU(cudaMemcpyAsync(back, c->d_scratch.p, sizeof(back), cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));   // the counts size the render launch; the caller's array is free again
    c->upload_recorded = false;             // complete: nothing for a caller's stream to wait for
    c->p_extent = reinterpret_cast<const unsigned int*>(c->d_scratch.p) + 2;
    c->p_box_fast = c->d_box_fast.p; c->p_box_strict = nullptr; c->p_box_cull = c->d_box_cull.p;
    c->p_obj_lrec = c->d_obj_lrec.p; c->p_sph = c->d_sph.p; c->p_box_obj = c->d_box_obj.p; c->p_sph_obj = c->d_sph_obj.p;
    c->p_cold = c->d_cold.p; c->p_lights = c->d_lights.p;
    c->n_box = (int)back[0];
    c->n_sph = (int)back[1];
    c->n_objects_dev = n;
    c->n_lights = (int)hs.lights.size();
    c->n_lit = hs.n_lit;
    c->n_lit_total = hs.n_lit_total;
    memcpy(c->lit, hs.lit, sizeof(c->lit)); memcpy(c->amb_sum, hs.amb_sum, sizeof(c->amb_sum));
    for (int k = 0; k < 3; k++) {
        const bool on = k < hs.n_lit_total && hs.n_lit_total <= 3;
        const RtLightDev z = {{0, 0, 0, 0}, {0, 0, 0, 0}, {0, 0, 0, 0}, {0, 0, 0, 0}};
        const RtLightDev& L = on ? hs.lights[hs.lit[k]] : z;
        c->lit_pos[k] = make_float4(L.position[0], L.position[1], L.position[2], 0.0f);
        c->lit_dif[k] = make_float4(L.diffuse[0], L.diffuse[1], L.diffuse[2], L.diffuse[3]);
        c->lit_spc[k] = make_float4(L.specular[0], L.specular[1], L.specular[2], L.specular[3]);
    }
    c->max_depth = hs.max_depth;
    c->have_scene = true;
    c->strict_ready = false;
    c->room_skip = 0;
    {
        const int rc2 = build_groups(c);
        if (rc2 != RT_OK) return rc2;
        if (c->n_groups > 0) { CU(cudaEventRecord(c->ev_upload, c->stream)); c->upload_recorded = true; }
    }
    return RT_OK;
}

int rt_set_prologue(rt_context* c, int mode) {
    if (!c) return fail(RT_ERR_BAD_ARG, "ctx is NULL");
    if (mode < RT_PROLOGUE_AUTO || mode > RT_PROLOGUE_DEVICE) return fail(RT_ERR_BAD_ARG, "bad prologue mode %d", mode);
    c->prologue_mode = mode;
    return RT_OK;
}

int rt_upload_scene(rt_context* c, const rt_scene* scene) {
    if (!c) return fail(RT_ERR_BAD_ARG, "ctx is NULL");
    int rc = validate_scene(scene);
    if (rc != RT_OK) return rc;
    const bool dev = c->prologue_mode == RT_PROLOGUE_DEVICE ||
                     (c->prologue_mode == RT_PROLOGUE_AUTO && scene->objects_count >= RT_PROLOGUE_DEVICE_MIN_OBJECTS);
    if (dev) {
        // a strict render after a device prologue needs the host records: the lights are kept, the objects stay
        // on the device (d_objs) and are fetched back only then
        c->lights_copy.assign(scene->lights, scene->lights + (scene->lights_count > 0 ? scene->lights_count : 0));
        c->depth_copy = scene->max_depth;
    }
    return dev ? upload_device(c, scene) : upload_host(c, scene);
}

int rt_render_uploaded(rt_context* c, const rt_camera* cameras, int n_frames,
                       void* out_rgba, int width, int height, const rt_options* opts_in) {
    if (!c) return fail(RT_ERR_BAD_ARG, "ctx is NULL");
    if (!c->have_scene) return fail(RT_ERR_BAD_ARG, "no scene uploaded");
    if (opts_in && opts_in->strict && !c->strict_ready) {
        // the scene went through the device prologue: the strict kernel wants the host run's records
        c->objs_copy.resize(c->n_objects_dev);
        CU(cudaSetDevice(c->device));
        CU(cudaStreamSynchronize(c->stream));
        if (c->n_objects_dev)
            CU(cudaMemcpy(c->objs_copy.data(), c->d_objs.p, c->n_objects_dev * sizeof(rt_object), cudaMemcpyDeviceToHost));
        rt_scene sc;
        sc.objects = c->objs_copy.data(); sc.objects_count = (int32_t)c->objs_copy.size();
        sc.lights = c->lights_copy.data(); sc.lights_count = (int32_t)c->lights_copy.size();
        sc.max_depth = c->depth_copy;
        int rc = upload_host(c, &sc);
        if (rc != RT_OK) return rc;
    }
    if (!cameras || n_frames <= 0) return fail(RT_ERR_BAD_ARG, "bad cameras / n_frames");
    if (!out_rgba) return fail(RT_ERR_BAD_ARG, "out_rgba is NULL");
    if (width <= 0 || height <= 0) return fail(RT_ERR_BAD_ARG, "bad image size %dx%d", width, height);
    rt_options od;
    rt_options_default(&od);
    const rt_options& o = opts_in ? *opts_in : od;
    const int spp = o.spp > 0 ? o.spp : 1;
    const int y0 = o.y0 > 0 ? o.y0 : 0;
    const int y1 = (o.y1 > 0 && o.y1 <= height) ? o.y1 : height;
    if (y0 >= y1) return fail(RT_ERR_BAD_ARG, "empty row band [%d,%d)", y0, y1);
    if (o.async && !o.out_is_device) return fail(RT_ERR_BAD_ARG, "async requires device output");
    if (o.device >= 0 && o.device != c->device)
        return fail(RT_ERR_BAD_ARG, "options.device %d != context device %d", o.device, c->device);
    CU(cudaSetDevice(c->device));
    cudaStream_t st = o.stream ? (cudaStream_t)o.stream : c->stream;
    // a caller's stream is not ordered with the context's: the scene upload must have landed before this render reads it
    if (st != c->stream && c->upload_recorded) CU(cudaStreamWaitEvent(st, c->ev_upload, 0));

    // camera prologue (glsl:366-383 hoisted)
    std::vector<RtCameraDev> cams(n_frames);
    for (int k = 0; k < n_frames; k++) {
        const rt_camera& cam = cameras[k];
        rthm::inverse_proj_view(cam.position, cam.angles, cam.v_fov, cam.aspect, cam.near_plane, cam.far_plane,
                                cams[k].inv_pv);
        cams[k].position[0] = cam.position[0]; cams[k].position[1] = cam.position[1];
        cams[k].position[2] = cam.position[2]; cams[k].position[3] = 1.0f;
    }
    if (o.stats) CU(cudaEventRecord(c->ev[2], st));
    // one frame: the 80 bytes of camera uniforms ride in the kernel parameters (no copy); a batch goes up as an array
    const bool cam_inline = n_frames == 1;
    if (!cam_inline) {
        CU(c->d_cams.reserve(n_frames));
        CU(cudaMemcpyAsync(c->d_cams.p, cams.data(), sizeof(RtCameraDev) * n_frames, cudaMemcpyHostToDevice, st));
    }

    const int band_rows = y1 - y0;
    const bool band_only = o.out_band_only != 0;
    const size_t frame_bytes_full = (size_t)width * height * 4;
    const size_t frame_bytes_band = (size_t)width * band_rows * 4;

    RtParams P;
    memset(&P, 0, sizeof(P));
    P.box_hot = reinterpret_cast<const float4*>(o.strict ? c->p_box_strict : c->p_box_fast);
    P.box_cull = reinterpret_cast<const float4*>(c->p_box_cull);
    P.sph_hot = reinterpret_cast<const float4*>(c->p_sph);
    P.obj_lrec = reinterpret_cast<const float4*>(c->p_obj_lrec);
    P.box_obj = c->p_box_obj; P.sph_obj = c->p_sph_obj;
    P.cold = c->p_cold; P.lights = c->p_lights; P.cams = cam_inline ? nullptr : c->d_cams.p;
    if (cam_inline) P.cam0 = cams[0];
    P.counters = c->d_cnt.p;
    P.tile_counter = c->d_tile.p;
    P.n_box = c->n_box; P.n_sph = c->n_sph; P.n_lights = c->n_lights; P.max_depth = c->max_depth;
    P.n_frames = n_frames; P.width = width; P.height = height; P.y0 = y0; P.y1 = y1;
    P.spp = spp; P.seed = o.seed;
    int n = 1;
    while ((n + 1) * (n + 1) <= spp) n++;
    P.spp_n = (n * n == spp) ? n : 0;
    P.tiles_x = (width + 7) / 8; P.tiles_y = (band_rows + 3) / 4;
    // Large scenes (fast kernel): the slab test runs behind a conservative bounding-sphere test, and
    // only the 16-byte bounding-sphere records are staged / streamed (rt_kernel.cuh: boxes_tile_cull).
    P.cull = (!o.strict && !(o.no_cull & 1)) ? 1 : 0;
    // Packet (cone) tests in front of the per-ray bounding tests; the per-light records are then only a
    // fallback for incoherent packets and are left out (they would halve the resident capacity).
    P.packet = (P.cull && !(o.no_cull & 4) && c->n_box + c->n_sph >= RT_PACKET_MIN_OBJECTS) ? 1 : 0;
    const int n_hot_all = c->n_box + c->n_sph;
    const bool block_pays = n_hot_all >= RT_BLOCK_MIN_OBJECTS && (c->max_depth == 0 || n_hot_all < RT_BLOCK_MAX_OBJECTS_BOUNCE);
    P.block = (P.packet && !(o.no_cull & 8) && (block_pays || (o.no_cull & 16))) ? 1 : 0;
    P.fan = (P.block && !(o.no_cull & 32)) ? 1 : 0;
    // bounce rays as grouped packets: group radius 0.1 E, directions within 8 degrees of the leader's (measured flat
    // between 0.02 E and 0.14 E and between 2.5 and 25 degrees on the 1k and 16k scenes)
    {
        // tuning knobs, read once: RT_GROUP_K (group radius^2 / E^2), RT_GROUP_COS (direction spread of a group)
        static const float group_k = getenv("RT_GROUP_K") ? (float)atof(getenv("RT_GROUP_K")) : 0.01f;
        static const float group_cos = getenv("RT_GROUP_COS") ? (float)atof(getenv("RT_GROUP_COS")) : 0.94f;
        P.group_k = (P.packet && !(o.no_cull & 64)) ? group_k : 0.0f;
        P.group_cos = group_cos;
    }
    P.extent = reinterpret_cast<const float*>(c->p_extent);
    P.room_skip = (c->room_skip && !o.strict && !(o.no_cull & 128)) ? 1 : 0;     // bit 7 (test switch): no inside-box shadow skip
    P.split_cones = (o.no_cull & 512) ? 0 : 1;                                   // bit 9 (test switch): one shadow cone per warp and light
    // groups of neighbouring objects: with bounce rays they pay from about 4k objects (depth 2, 1080p: 2k objects 10.3
    // against 9.9 ms without, 4k 15.6 / 15.9, 16k 30.6 / 39.1; depth 0: 4k 0.47 / 0.51, 16k 0.77 / 0.93, 64k 1.15 / 2.01)
    const bool groups_pay = c->max_depth == 0 || c->n_box + c->n_sph >= RT_GROUPS_MIN_OBJECTS_BOUNCE;
    if (P.packet && c->n_groups > 0 && groups_pay && !(o.no_cull & 256)) {       // bit 8 (test switch): packet tests object by object
        P.grp = reinterpret_cast<const float4*>(c->d_grp.p); P.srec = reinterpret_cast<const float4*>(c->d_srec.p);
        P.sidx = c->d_sidx.p; P.n_groups = c->n_groups;
    }
    if (P.block) {
        // one list slice per CTA of the persistent grid (at most RT_MAX_CTAS_PER_SM per SM); every hot object fits
        P.blist_stride = c->n_box + c->n_sph;
        CU(c->d_blist.reserve((size_t)RT_MAX_CTAS_PER_SM * c->sm_count * (size_t)P.blist_stride));
        P.blist = c->d_blist.p;
        P.brec_cap = P.blist_stride < RT_BLOCK_REC_CAP ? P.blist_stride : RT_BLOCK_REC_CAP;
        CU(c->d_brec.reserve((size_t)RT_MAX_CTAS_PER_SM * c->sm_count * (size_t)P.brec_cap * 4));
        P.brec = reinterpret_cast<float4*>(c->d_brec.p);
    }
    P.n_lit = (P.cull && !P.packet && !(o.no_cull & 2)) ? c->n_lit : 0;
    P.n_lit_batch = c->n_lit_total <= 3 ? c->n_lit_total : -1;
    memcpy(P.lit, c->lit, sizeof(P.lit)); memcpy(P.amb_sum, c->amb_sum, sizeof(P.amb_sum));
    memcpy(P.lit_pos, c->lit_pos, sizeof(P.lit_pos)); memcpy(P.lit_dif, c->lit_dif, sizeof(P.lit_dif));
    memcpy(P.lit_spc, c->lit_spc, sizeof(P.lit_spc));
    size_t hot_bytes = ((size_t)P.n_box * (o.strict ? RT_BOX_F4_STRICT : (P.cull ? 1 : RT_BOX_F4_FAST)) + P.n_sph) * 16;
    if (P.n_lit > 0) hot_bytes += (size_t)((P.n_box + P.n_sph + 63) / 64) * P.n_lit * 64 * 16;
    P.resident = hot_bytes <= (o.strict ? rt_strict::resident_limit_bytes(c->max_smem_optin, false)
                                        : rt_fast::resident_limit_bytes(c->max_smem_optin, P.cull != 0)) ? 1 : 0;

    // Page-locked host output (rt_host_alloc): the kernel's 16-byte framebuffer stores can go straight
    // to host memory over PCIe (posted writes, overlapped with the rendering) instead of a device frame
    // plus a device->host copy after the kernel.
    uint8_t* host_mapped = nullptr;
    if (!o.out_is_device && o.host_store != RT_HOST_STORE_NEVER) {
        cudaPointerAttributes pa;
        if (cudaPointerGetAttributes(&pa, out_rgba) == cudaSuccess && pa.type == cudaMemoryTypeHost && pa.devicePointer) {
            // auto: worth it when the frame takes clearly longer to render than to push over PCIe
            // (bounce depth >= 1 or many objects or several samples); a depth-0 built-in frame is faster with the copy
            const bool heavy = c->max_depth >= 1 || c->n_box + c->n_sph >= 48 || spp > 1;
            if (o.host_store == RT_HOST_STORE_ALWAYS || heavy) host_mapped = (uint8_t*)pa.devicePointer;
        } else {
            (void)cudaGetLastError();   // pageable memory: not an error
        }
    }
    if (o.out_is_device || host_mapped) {
        P.out = o.out_is_device ? (uint8_t*)out_rgba : host_mapped;
        P.frame_stride = band_only ? (long long)frame_bytes_band : (long long)frame_bytes_full;
        P.out_row0 = band_only ? y0 : 0;
    } else {
        // render the band compactly into the context's device buffer, then copy out
        CU(c->d_out.reserve(frame_bytes_band * n_frames));
        P.out = c->d_out.p;
        P.frame_stride = (long long)frame_bytes_band;
        P.out_row0 = y0;
    }

    const int grid = c->sm_count;   // the launcher sizes the persistent grid from the occupancy of the chosen kernel

    const bool count = o.stats != nullptr;
    if (count) CU(cudaMemsetAsync(c->d_cnt.p, 0, sizeof(RtCounters), st));
    // profiling aid: RT_BLOCK_LOG=path makes the instrumented twin of the block kernel write one 64-byte
    // record per block (start, duration, list sizes, per-warp times), dumped to `path` after the render
    const char* blk_log_path = (count && P.block) ? getenv("RT_BLOCK_LOG") : nullptr;
    size_t blk_log_words = 0;
    if (blk_log_path && *blk_log_path) {
        blk_log_words = (size_t)((P.tiles_x + 5) / 6) * ((P.tiles_y + 3) / 4) * n_frames * 16;
        CU(c->d_blklog.reserve(blk_log_words));
        CU(cudaMemsetAsync(c->d_blklog.p, 0, blk_log_words * 4, st));
    }
    CU(cudaMemsetAsync(c->d_tile.p, 0, 2 * sizeof(unsigned long long), st));
    if (o.stats) CU(cudaEventRecord(c->ev[0], st));
    CU(o.strict ? rt_strict::launch(P, grid, false, st) : rt_fast::launch(P, grid, false, st));
    g_launches++;
    if (o.stats) CU(cudaEventRecord(c->ev[1], st));
    if (count) {
        // instrumented twin: same kernel with counters, run after the timed one (results identical)
        CU(cudaMemsetAsync(c->d_tile.p, 0, 2 * sizeof(unsigned long long), st));
        if (blk_log_words) P.blk_log = c->d_blklog.p;
        CU(o.strict ? rt_strict::launch(P, grid, true, st) : rt_fast::launch(P, grid, true, st));
        P.blk_log = nullptr;
        g_launches++;
    }

    // Completion word (multi-GPU frame mode): raised in stream order behind the render kernel -- at a kernel
    // boundary every framebuffer store of the launch is visible, the one-thread kernel then stores the value with
    // system-scope release.  (Written by the render kernel's own last CTA it cost the small-scene kernel 4 % more
    // instructions and 6 % of its time: the epilogue upset the register allocation of the whole kernel.)
    if (o.done_flag) {
        if (rt_write_flag(reinterpret_cast<uint32_t*>(o.done_flag), o.done_value, st) != RT_OK)
            return fail(RT_ERR_CUDA, "completion-flag launch failed: %s", cudaGetErrorString(cudaGetLastError()));
        g_launches++;
    }
    if (!o.out_is_device && !host_mapped) {
        uint8_t* dst = (uint8_t*)out_rgba;
        for (int k = 0; k < n_frames; k++) {
            uint8_t* d = band_only ? dst + (size_t)k * frame_bytes_band
                                   : dst + (size_t)k * frame_bytes_full + (size_t)y0 * width * 4;
            CU(cudaMemcpyAsync(d, c->d_out.p + (size_t)k * frame_bytes_band, frame_bytes_band,
                               cudaMemcpyDeviceToHost, st));
        }
    }
    if (o.stats) CU(cudaEventRecord(c->ev[3], st));
    if (st != c->stream) {                  // the next upload must not overwrite what this render still reads
        CU(cudaEventRecord(c->ev_render, st));
        c->render_recorded = true;
    }
    if (!o.async) {
        CU(cudaStreamSynchronize(st));   // glFinish() (main.cpp:238)
    }
    if (o.stats) {
        if (o.async) CU(cudaStreamSynchronize(st));
        RtCounters hc;
        CU(cudaMemcpy(&hc, c->d_cnt.p, sizeof(hc), cudaMemcpyDeviceToHost));
        rt_stats* s = o.stats;
        s->rays = hc.rays; s->nodes = hc.nodes; s->hit_nodes = hc.hit_nodes; s->shadow_rays = hc.shadow_rays;
        s->box_tests = hc.box_tests; s->sphere_tests = hc.sphere_tests; s->red_pixels = hc.red_pixels;
        s->rays_not_traversed = hc.pad;
        s->cull_tests = hc.cull_tests; s->warp_iters = hc.warp_iters; s->packet_tests = hc.packet_tests;
        s->group_passes = hc.group_passes; s->loop_passes = hc.loop_passes;
        s->candidates = hc.candidates; s->shadow_group_passes = hc.shadow_group_passes;
        s->shaded_lights = hc.shaded_lights;
        CU(cudaEventElapsedTime(&s->kernel_ms, c->ev[0], c->ev[1]));
        CU(cudaEventElapsedTime(&s->total_ms, c->ev[2], c->ev[3]));
        if (blk_log_words) {
            std::vector<uint32_t> hl(blk_log_words);
            CU(cudaMemcpy(hl.data(), c->d_blklog.p, blk_log_words * 4, cudaMemcpyDeviceToHost));
            if (FILE* f = fopen(blk_log_path, "wb")) { fwrite(hl.data(), 4, blk_log_words, f); fclose(f); }
        }
    }
    return RT_OK;
}

// No counterpart in the reference (see include/rt_b200.h).
// A separate function on purpose: the existing render path is untouched.  Nothing below allocates -- no DevBuf
// reserve, no std::vector, no host staging: the camera records travel in the set-up kernel's parameters and all
// device scratch is the file-scope array of rt_kernel_fast.cu.
int rt_render_block_noalloc(rt_context* c, const rt_camera* cameras, int n_frames, void* out_rgba_device,
                            int width, int height, const rt_options* opts_in) {
    if (!c) return fail(RT_ERR_BAD_ARG, "ctx is NULL");
    if (!c->have_scene) return fail(RT_ERR_BAD_ARG, "no scene uploaded (rt_upload_scene first: the no-alloc call does no set-up)");
    if (!cameras || n_frames <= 0) return fail(RT_ERR_BAD_ARG, "bad cameras / n_frames");
    if (!out_rgba_device) return fail(RT_ERR_BAD_ARG, "out_rgba_device is NULL");
    if (width <= 0 || height <= 0) return fail(RT_ERR_BAD_ARG, "bad image size %dx%d", width, height);
    rt_options od;
    rt_options_default(&od);
    const rt_options& o = opts_in ? *opts_in : od;
    if (o.strict) return fail(RT_ERR_UNSUPPORTED, "rt_render_block_noalloc: the strict kernel is not available here");
    if (o.stats) return fail(RT_ERR_UNSUPPORTED, "rt_render_block_noalloc: no counters (rt_options.stats must be NULL)");
    if (o.no_cull & (1 | 4 | 8)) return fail(RT_ERR_UNSUPPORTED, "rt_render_block_noalloc: no_cull bits 0, 2, 3 switch the block kernel off");
    const int n_hot = c->n_box + c->n_sph;
    if (n_frames > RT_NOALLOC_MAX_FRAMES)
        return fail(RT_ERR_UNSUPPORTED, "rt_render_block_noalloc: %d cameras, the static scratch holds at most RT_NOALLOC_MAX_FRAMES = %d",
                    n_frames, RT_NOALLOC_MAX_FRAMES);
    if (n_hot > RT_NOALLOC_MAX_OBJECTS)
        return fail(RT_ERR_UNSUPPORTED, "rt_render_block_noalloc: %d boxes + spheres, the static scratch holds the candidate lists of at most "
                    "RT_NOALLOC_MAX_OBJECTS = %d (use rt_render_uploaded)", n_hot, RT_NOALLOC_MAX_OBJECTS);
    if (n_hot < RT_PACKET_MIN_OBJECTS)
        return fail(RT_ERR_UNSUPPORTED, "rt_render_block_noalloc: %d boxes + spheres, the block kernel needs at least %d (use rt_render_uploaded)",
                    n_hot, RT_PACKET_MIN_OBJECTS);
    const int spp = o.spp > 0 ? o.spp : 1;
    const int y0 = o.y0 > 0 ? o.y0 : 0;
    const int y1 = (o.y1 > 0 && o.y1 <= height) ? o.y1 : height;
    if (y0 >= y1) return fail(RT_ERR_BAD_ARG, "empty row band [%d,%d)", y0, y1);
    if (o.device >= 0 && o.device != c->device)
        return fail(RT_ERR_BAD_ARG, "options.device %d != context device %d", o.device, c->device);
    CU(cudaSetDevice(c->device));
    cudaStream_t st = o.stream ? (cudaStream_t)o.stream : c->stream;
    if (st != c->stream && c->upload_recorded) CU(cudaStreamWaitEvent(st, c->ev_upload, 0));

    RtCameraDev cams[RT_NOALLOC_MAX_FRAMES];            // camera prologue (glsl:366-383 hoisted), on the stack
    for (int k = 0; k < n_frames; k++) {
        const rt_camera& cam = cameras[k];
        rthm::inverse_proj_view(cam.position, cam.angles, cam.v_fov, cam.aspect, cam.near_plane, cam.far_plane, cams[k].inv_pv);
        cams[k].position[0] = cam.position[0]; cams[k].position[1] = cam.position[1];
        cams[k].position[2] = cam.position[2]; cams[k].position[3] = 1.0f;
    }
    const int band_rows = y1 - y0;
    const bool band_only = o.out_band_only != 0;
    RtParams P;
    memset(&P, 0, sizeof(P));
    P.box_hot = reinterpret_cast<const float4*>(c->p_box_fast);
    P.box_cull = reinterpret_cast<const float4*>(c->p_box_cull);
    P.sph_hot = reinterpret_cast<const float4*>(c->p_sph);
    P.obj_lrec = reinterpret_cast<const float4*>(c->p_obj_lrec);
    P.box_obj = c->p_box_obj; P.sph_obj = c->p_sph_obj;
    P.cold = c->p_cold; P.lights = c->p_lights;
    P.counters = c->d_cnt.p;
    P.n_box = c->n_box; P.n_sph = c->n_sph; P.n_lights = c->n_lights; P.max_depth = c->max_depth;
    P.n_frames = n_frames; P.width = width; P.height = height; P.y0 = y0; P.y1 = y1;
    P.spp = spp; P.seed = o.seed;
    int n = 1;
    while ((n + 1) * (n + 1) <= spp) n++;
    P.spp_n = (n * n == spp) ? n : 0;
    P.tiles_x = (width + 7) / 8; P.tiles_y = (band_rows + 3) / 4;
    P.cull = 1; P.packet = 1; P.block = 1;              // this entry point IS the block kernel
    P.fan = (o.no_cull & 32) ? 0 : 1;
    P.split_cones = (o.no_cull & 512) ? 0 : 1;
    {
        static const float group_k = getenv("RT_GROUP_K") ? (float)atof(getenv("RT_GROUP_K")) : 0.01f;
        static const float group_cos = getenv("RT_GROUP_COS") ? (float)atof(getenv("RT_GROUP_COS")) : 0.94f;
        P.group_k = (o.no_cull & 64) ? 0.0f : group_k;
        P.group_cos = group_cos;
    }
    P.extent = reinterpret_cast<const float*>(c->p_extent);
    P.room_skip = 0;                                    // small-scene kernels only
    P.n_lit = 0;                                        // packet tests: no per-light records
    P.n_lit_batch = c->n_lit_total <= 3 ? c->n_lit_total : -1;
    memcpy(P.lit, c->lit, sizeof(P.lit)); memcpy(P.amb_sum, c->amb_sum, sizeof(P.amb_sum));
    memcpy(P.lit_pos, c->lit_pos, sizeof(P.lit_pos)); memcpy(P.lit_dif, c->lit_dif, sizeof(P.lit_dif));
    memcpy(P.lit_spc, c->lit_spc, sizeof(P.lit_spc));
    const size_t hot_bytes = ((size_t)P.n_box + P.n_sph) * 16;
    P.resident = hot_bytes <= rt_fast::resident_limit_bytes(c->max_smem_optin, true) ? 1 : 0;
    P.out = (uint8_t*)out_rgba_device;
    P.frame_stride = band_only ? (long long)width * band_rows * 4 : (long long)width * height * 4;
    P.out_row0 = band_only ? y0 : 0;

    const char* why = "";
    cudaError_t e = rt_fast::launch_block_noalloc(P, cams, c->device, c->sm_count, st, &why);
    if (e == cudaErrorInvalidValue && why[0]) return fail(RT_ERR_UNSUPPORTED, "rt_render_block_noalloc: %s", why);
    if (e != cudaSuccess) return fail(RT_ERR_CUDA, "rt_render_block_noalloc: %s (%s)", cudaGetErrorString(e), why);
    g_launches += 2;                                    // set-up kernel + render kernel
    if (o.done_flag) {
        if (rt_write_flag(reinterpret_cast<uint32_t*>(o.done_flag), o.done_value, st) != RT_OK)
            return fail(RT_ERR_CUDA, "completion-flag launch failed: %s", cudaGetErrorString(cudaGetLastError()));
        g_launches++;
    }
    if (st != c->stream) {
        CU(cudaEventRecord(c->ev_render, st));
        c->render_recorded = true;
    }
    if (!o.async) CU(cudaStreamSynchronize(st));
    return RT_OK;
}

int rt_render_batch(rt_context* c, const rt_scene* scene, const rt_camera* cameras, int n_frames,
                    void* out_rgba, int width, int height, const rt_options* opts) {
    int rc = rt_upload_scene(c, scene);
    if (rc != RT_OK) return rc;
    return rt_render_uploaded(c, cameras, n_frames, out_rgba, width, height, opts);
}

int rt_render_ex(rt_context* c, const rt_scene* scene, const rt_camera* camera, void* out_rgba,
                 int width, int height, const rt_options* opts) {
    if (!camera) return fail(RT_ERR_BAD_ARG, "camera_uniforms is NULL");
    return rt_render_batch(c, scene, camera, 1, out_rgba, width, height, opts);
}

int rt_render(const rt_scene* scene, const rt_camera* camera, uint8_t* out_rgba, int width, int height) {
    static std::mutex mu;
    static rt_context* ctxs[64] = {nullptr};
    int dev = 0;
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0)
        return fail(RT_ERR_NO_DEVICE, "no CUDA device (%s); this library has no CPU fallback",
                    e != cudaSuccess ? cudaGetErrorString(e) : "device count 0");
    CU(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) return fail(RT_ERR_BAD_ARG, "device ordinal %d unsupported", dev);
    std::lock_guard<std::mutex> lock(mu);
    if (!ctxs[dev]) {
        int rc = rt_create(&ctxs[dev], dev);
        if (rc != RT_OK) return rc;
    }
    return rt_render_ex(ctxs[dev], scene, camera, out_rgba, width, height, nullptr);
}

int rt_device_alloc(void** device_ptr, size_t bytes) {
    if (!device_ptr || bytes == 0) return fail(RT_ERR_BAD_ARG, "bad rt_device_alloc arguments");
    CU(cudaMalloc(device_ptr, bytes));
    return RT_OK;
}
int rt_device_free(void* device_ptr) {
    if (device_ptr) CU(cudaFree(device_ptr));
    return RT_OK;
}

int rt_ipc_export(void* device_ptr, uint8_t handle64[64]) {
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
    if (!device_ptr || !handle64) return fail(RT_ERR_BAD_ARG, "bad rt_ipc_export arguments");
    cudaIpcMemHandle_t h;
    CU(cudaIpcGetMemHandle(&h, device_ptr));
    memcpy(handle64, &h, 64);
    return RT_OK;
}
int rt_ipc_import(const uint8_t handle64[64], void** device_ptr) {
    if (!device_ptr || !handle64) return fail(RT_ERR_BAD_ARG, "bad rt_ipc_import arguments");
    cudaIpcMemHandle_t h;
    memcpy(&h, handle64, 64);
    CU(cudaIpcOpenMemHandle(device_ptr, h, cudaIpcMemLazyEnablePeerAccess));
    return RT_OK;
}
int rt_ipc_close(void* device_ptr) {
    if (device_ptr) CU(cudaIpcCloseMemHandle(device_ptr));
    return RT_OK;
}

}  // extern "C"
