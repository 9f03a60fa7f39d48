/*
 * rt_prologue.cu — the host prologue as device kernels.
 *
 * The shader rebuilds local_to_world, its inverse and the normal matrix for every (ray, box) test
 * (glsl:650-652, 718); the library hoists that to once per object.  For large scenes re-uploaded every
 * frame (an animated 64k-object scene: 64k x sin/cos/inverse on one host thread + 27 MB of records
 * over PCIe) the hoisted work itself becomes the frame's bottleneck, so it moves to the GPU: the raw
 * 132-byte objects go up in ONE copy straight from the caller's array (8.6 MB at 64k; at full PCIe rate when
 * the array is page-locked, rt_host_alloc) and ONE kernel classifies them (glsl:749, 759), assigns the
 * hot-list positions (a scan that keeps the original order, glsl:744) and builds every record in HBM.  The
 * host never touches the objects; it reads back three words (box count, sphere count, scene extent).
 * Same functions as the host run (rt_prologue_math.h), compiled -fmad=false; only sinf / cosf differ
 * (CUDA vs glibc), so the strict kernel keeps the host prologue.
 */
#include <cuda_runtime.h>
#include <stdint.h>

#include "rt_prologue_math.h"

namespace rt_prologue {

constexpr int kThreads = 256;

// Grid-wide barrier of a grid whose CTAs are all resident (the launcher sizes it from the occupancy API and the
// kernel strides over its work): one arrival counter per barrier, zeroed by the launcher.
__device__ __forceinline__ void grid_barrier(unsigned int* counter) {
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        atomicAdd(counter, 1u);
        while (atomicAdd(counter, 0u) < gridDim.x) __nanosleep(32);
        __threadfence();
    }
    __syncthreads();
}

/* out[0] = n_box, out[1] = n_sph, out[2] = scene extent bits; sync[0..1] = barrier counters */
struct Scratch {
    unsigned int n_box, n_sph, extent_bits, pad_;
    unsigned int sync[4];
};

// CTA b owns the objects [b * chunk, (b + 1) * chunk).
//   phase 0  kind of every object (glsl:749, 759) and the CTA's box / sphere counts
//   phase 1  hot-list positions (exclusive scan: counts of the CTAs before this one + ballot scan inside the chunk,
//            original order kept, glsl:744), cold + hot records of every object, scene extent (max)
//   phase 2  what needs the final extent: padded bounding radii, per-light records
__global__ void __launch_bounds__(kThreads) fused_kernel(const rt_object* __restrict__ objs, int n, int chunk,
                                                         unsigned int* __restrict__ cta_counts,      // [gridDim.x][2]
                                                         int32_t* __restrict__ box_obj, int32_t* __restrict__ sph_obj,
                                                         RtCold* __restrict__ cold, float* __restrict__ box_fast,
                                                         float* __restrict__ box_cull, float* __restrict__ sph,
                                                         const RtLightDev* __restrict__ lights, int n_lit, int lit0, int lit1, int lit2,
                                                         float* __restrict__ obj_lrec, Scratch* __restrict__ sc) {
    __shared__ unsigned int s_warp[kThreads / 32][2];
    __shared__ unsigned int s_base[2];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int first = blockIdx.x * chunk, last = min(n, first + chunk);

    // ---- phase 0: counts
    unsigned int nb = 0, ns = 0;
    for (int i = first + threadIdx.x; i < last; i += kThreads) {
        const int k = rtpm::object_kind(objs[i]);
        nb += k == RT_KIND_BOX; ns += k == RT_KIND_SPHERE;
    }
    for (int off = 16; off > 0; off >>= 1) { nb += __shfl_down_sync(0xffffffffu, nb, off); ns += __shfl_down_sync(0xffffffffu, ns, off); }
    if (lane == 0) { s_warp[warp][0] = nb; s_warp[warp][1] = ns; }
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned int b = 0, s2 = 0;
        for (int k = 0; k < kThreads / 32; k++) { b += s_warp[k][0]; s2 += s_warp[k][1]; }
        cta_counts[2 * blockIdx.x] = b; cta_counts[2 * blockIdx.x + 1] = s2;
    }
    grid_barrier(&sc->sync[0]);

    // ---- phase 1: positions and records
    {
        unsigned int pb = 0, ps = 0, tb = 0, ts = 0;
        for (int b = threadIdx.x; b < (int)gridDim.x; b += kThreads) {
            const unsigned int cb = __ldcg(cta_counts + 2 * b), cs = __ldcg(cta_counts + 2 * b + 1);
            tb += cb; ts += cs;
            if (b < (int)blockIdx.x) { pb += cb; ps += cs; }
        }
        for (int off = 16; off > 0; off >>= 1) {
            pb += __shfl_down_sync(0xffffffffu, pb, off); ps += __shfl_down_sync(0xffffffffu, ps, off);
            tb += __shfl_down_sync(0xffffffffu, tb, off); ts += __shfl_down_sync(0xffffffffu, ts, off);
        }
        __syncthreads();
        if (lane == 0) { s_warp[warp][0] = pb; s_warp[warp][1] = ps; }
        __shared__ unsigned int s_tot[kThreads / 32][2];
        if (lane == 0) { s_tot[warp][0] = tb; s_tot[warp][1] = ts; }
        __syncthreads();
        if (threadIdx.x == 0) {
            unsigned int b = 0, s2 = 0, b2 = 0, s3 = 0;
            for (int k = 0; k < kThreads / 32; k++) { b += s_warp[k][0]; s2 += s_warp[k][1]; b2 += s_tot[k][0]; s3 += s_tot[k][1]; }
            s_base[0] = b; s_base[1] = s2;
            if (blockIdx.x == 0) { sc->n_box = b2; sc->n_sph = s3; }
        }
        __syncthreads();
    }
    float reach = 0.0f;
    for (int i0 = first; i0 < last; i0 += kThreads) {            // CTA-uniform trip count
        const int i = i0 + threadIdx.x;
        int k = RT_KIND_SKIP;
        rt_object o;
        if (i < last) { o = objs[i]; k = rtpm::object_kind(o); }
        // in-chunk exclusive scan in original order: ballot inside the warp, warp totals through shared memory
        const unsigned int mb = __ballot_sync(0xffffffffu, k == RT_KIND_BOX), ms = __ballot_sync(0xffffffffu, k == RT_KIND_SPHERE);
        const unsigned int lt = (1u << lane) - 1u;
        __syncthreads();
        if (lane == 0) { s_warp[warp][0] = __popc(mb); s_warp[warp][1] = __popc(ms); }
        __syncthreads();
        unsigned int wb = 0, ws = 0, allb = 0, alls = 0;
        for (int q = 0; q < kThreads / 32; q++) {
            if (q < warp) { wb += s_warp[q][0]; ws += s_warp[q][1]; }
            allb += s_warp[q][0]; alls += s_warp[q][1];
        }
        const int hb = (int)(s_base[0] + wb + __popc(mb & lt)), hs = (int)(s_base[1] + ws + __popc(ms & lt));
        if (i < last) {
            const int h = k == RT_KIND_BOX ? hb : (k == RT_KIND_SPHERE ? hs : -1);
            RtCold c;
            float fast12[12], cull4[4], sph4[4];
            rtpm::object_records(o, k, h, &c, fast12, nullptr, cull4, sph4);
            cold[i] = c;
            if (k == RT_KIND_BOX) {
                for (int q = 0; q < 12; q++) box_fast[(size_t)h * 12 + q] = fast12[q];
                for (int q = 0; q < 4; q++) box_cull[(size_t)h * 4 + q] = cull4[q];     // radius still un-padded
                box_obj[h] = i;
                reach = fmaxf(reach, (float)rtpm::reach(cull4, fabs((double)cull4[3])));
            } else if (k == RT_KIND_SPHERE) {
                for (int q = 0; q < 4; q++) sph[(size_t)h * 4 + q] = sph4[q];
                sph_obj[h] = i;
                reach = fmaxf(reach, (float)rtpm::reach(sph4, sqrt(fabs((double)sph4[3]))));
            }
        }
        __syncthreads();
        if (threadIdx.x == 0) { s_base[0] += allb; s_base[1] += alls; }
        __syncthreads();
    }
    // scene extent: max over objects (non-negative floats order like their bit patterns)
    for (int off = 16; off > 0; off >>= 1) reach = fmaxf(reach, __shfl_down_sync(0xffffffffu, reach, off));
    if (lane == 0 && reach > 0.0f) atomicMax(&sc->extent_bits, __float_as_uint(reach * 1.000001f));
    grid_barrier(&sc->sync[1]);

    // ---- phase 2: per hot object (boxes then spheres): per-light records, then the padded bounding radius
    const int n_box = (int)__ldcg(&sc->n_box), n_sph = (int)__ldcg(&sc->n_sph);
    const double extent = (double)__uint_as_float(__ldcg(&sc->extent_bits));
    const int n_hot = n_box + n_sph;
    for (int h = blockIdx.x * kThreads + threadIdx.x; h < n_hot; h += gridDim.x * kThreads) {
        const bool is_b = h < n_box;
        float c4[4];
        for (int q = 0; q < 4; q++) c4[q] = is_b ? __ldcg(box_cull + (size_t)h * 4 + q) : __ldcg(sph + (size_t)(h - n_box) * 4 + q);
        const double rad = is_b ? (double)c4[3] : sqrt((double)c4[3]);
        for (int s2 = 0; s2 < n_lit; s2++) {
            const int li = s2 == 0 ? lit0 : (s2 == 1 ? lit1 : lit2);
            float rec[4];
            rtpm::light_record(c4, rad, is_b, lights[li].position, extent, rec);
            float* dst = obj_lrec + (((size_t)(h / 64) * n_lit + s2) * 64 + (h % 64)) * 4;
            for (int q = 0; q < 4; q++) dst[q] = rec[q];
        }
        if (is_b) box_cull[(size_t)h * 4 + 3] = rtpm::padded_radius2((double)c4[3], extent);
    }
    if (n_lit > 0) {     // the padding of the last light-record tile: records that never pass
        const int n_f4 = ((n_hot + 63) / 64) * 64;
        for (int h = n_hot + blockIdx.x * kThreads + threadIdx.x; h < n_f4; h += gridDim.x * kThreads)
            for (int s2 = 0; s2 < n_lit; s2++)
                reinterpret_cast<float4*>(obj_lrec)[((size_t)(h / 64) * n_lit + s2) * 64 + (h % 64)] = make_float4(0.0f, 0.0f, 0.0f, 1e30f);
    }
}

int max_grid(int sm_count);
size_t scratch_bytes(int max_ctas) { return sizeof(Scratch) + (size_t)max_ctas * 2 * sizeof(unsigned int); }

// d_scratch: scratch_bytes(grid) bytes.  After the kernel, d_scratch holds {n_box, n_sph, extent bits}.
cudaError_t run(const rt_object* d_objs, int n_obj, int32_t* d_box_obj, int32_t* d_sph_obj, RtCold* d_cold,
                float* d_box_fast, float* d_box_cull, float* d_sph, const RtLightDev* d_lights, int n_lit, const int* lit,
                float light_reach, float* d_obj_lrec, void* d_scratch, int sm_count, cudaStream_t st) {
    // every CTA must be resident (grid barriers): at most one wave; chunks of whole 256-object rows
    int grid = (n_obj + kThreads - 1) / kThreads;
    const int cap = max_grid(sm_count);
    if (grid > cap) grid = cap;
    if (grid < 1) grid = 1;
    int chunk = (n_obj + grid - 1) / grid;
    chunk = (chunk + kThreads - 1) / kThreads * kThreads;
    Scratch init = {};
    init.extent_bits = __builtin_bit_cast(unsigned int, light_reach > 1.0f ? light_reach : 1.0f);
    cudaError_t e = cudaMemcpyAsync(d_scratch, &init, sizeof(init), cudaMemcpyHostToDevice, st);
    if (e != cudaSuccess) return e;
    Scratch* sc = reinterpret_cast<Scratch*>(d_scratch);
    fused_kernel<<<grid, kThreads, 0, st>>>(d_objs, n_obj, chunk, reinterpret_cast<unsigned int*>(sc + 1), d_box_obj, d_sph_obj,
                                            d_cold, d_box_fast, d_box_cull, d_sph, d_lights, n_lit, lit[0], lit[1], lit[2],
                                            d_obj_lrec, sc);
    return cudaGetLastError();
}

int max_grid(int sm_count) {
    int per_sm = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, fused_kernel, kThreads, 0) != cudaSuccess || per_sm < 1) per_sm = 1;
    if (per_sm > 2) per_sm = 2;          // plenty for a few hundred flops per object, and far inside what is co-resident
    return per_sm * sm_count;
}

}  // namespace rt_prologue
