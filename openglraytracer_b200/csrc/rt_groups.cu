/*
 * rt_groups.cu -- bounding spheres over groups of 32 neighbouring objects, for the packet tests of the large-scene
 * kernels (block_filter, packet_trace: DESIGN.md 4).
 *
 * The shader tests every ray against every object (glsl:744).  The large-scene kernels first ask, per packet of
 * rays, which objects' bounding spheres the packet's cone can meet at all; that question is asked here one level
 * up: the hot list (boxes, then spheres, in the objects' original order -- untouched) is read in Morton order of
 * the objects' centres, every 32 consecutive entries of that order form a group, and a group's sphere holds the
 * bounding spheres of its members.  A packet that misses a group's sphere misses every member (rt_cone.h:
 * group_pass), so its 32 member tests are skipped; the members of the other groups are tested exactly as before.
 * Nothing here changes which objects a ray is tested against in the end, nor the order-independent accept rule:
 * frames are byte-identical with the groups switched off (rt_options.no_cull bit 8).
 *
 * One kernel, all CTAs resident (grid barriers):
 *   0  bounding box of the centres
 *   1  24-bit Morton keys (8 bits per axis), stable LSD radix sort of (key, hot index) in three 8-bit passes --
 *      equal keys keep the hot-list order, so the result does not depend on scheduling
 *   2  one warp per group: the bounding records in sorted order (a copy: a group's 32 records are one 512-byte
 *      read), their hot-list positions, the group sphere
 */
#include <cuda_runtime.h>
#include <math_constants.h>
#include <stdint.h>

#include "rt_device.h"

namespace rt_groups {

constexpr int kThreads = 256;
constexpr int kWarps = kThreads / 32;
constexpr unsigned kFull = 0xffffffffu;

struct Scratch {
    unsigned int bb[6];          // ordered-int encodings: min x, y, z; max x, y, z
    unsigned int pad_[2];
    unsigned int sync[8];        // grid-barrier arrival counters
};

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
// floats as unsigned integers of the same order
__device__ __forceinline__ unsigned enc(float f) { const unsigned b = __float_as_uint(f); return (b & 0x80000000u) ? ~b : (b | 0x80000000u); }
__device__ __forceinline__ float dec(unsigned e) { return __uint_as_float((e & 0x80000000u) ? (e & 0x7fffffffu) : ~e); }
__device__ __forceinline__ unsigned spread8(unsigned v) {      // 8 bits -> every third bit
    v &= 0xffu;
    v = (v | (v << 8)) & 0x00f00fu;
    v = (v | (v << 4)) & 0x0c30c3u;
    v = (v | (v << 2)) & 0x249249u;
    return v;
}
__device__ __forceinline__ float sqrt_approx(float x) { float r; asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }

struct Arrays {
    const float4* box_cull; const float4* sph; int n_box, n_sph;
    const float* extent;
    unsigned int *key_a, *key_b, *val_a, *val_b;     // n_hot each
    unsigned int* hist;                              // [gridDim.x][256]
    float4* grp; float4* srec; uint32_t* sidx;       // outputs: n_groups, 32 n_groups, 32 n_groups
    Scratch* sc;
};

__device__ __forceinline__ float4 bounding_record(const Arrays& A, int hot) {
    return hot < A.n_box ? __ldg(A.box_cull + hot) : __ldg(A.sph + (hot - A.n_box));
}

// One stable pass over 8 bits of the key: CTA b owns the entries [first, last) of the source order.
__device__ void radix_pass(const Arrays& A, const unsigned* __restrict__ src_k, const unsigned* __restrict__ src_v,
                           unsigned* __restrict__ dst_k, unsigned* __restrict__ dst_v, int shift, int first, int last,
                           unsigned int* bar_h, unsigned int* bar_s) {
    __shared__ unsigned s_hist[256];
    __shared__ unsigned s_base[256];
    __shared__ unsigned s_wcount[kWarps][256];
    __shared__ unsigned s_scan[kWarps];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, t = threadIdx.x;
    // ---- histogram of this CTA's chunk
    s_hist[t] = 0u;
    __syncthreads();
    for (int i = first + t; i < last; i += kThreads) atomicAdd(&s_hist[(__ldcg(src_k + i) >> shift) & 255u], 1u);
    __syncthreads();
    A.hist[(size_t)blockIdx.x * 256 + t] = s_hist[t];
    grid_barrier(bar_h);
    // ---- where digit t of this CTA starts: all smaller digits, then digit t of the CTAs before this one
    unsigned before = 0u, total = 0u;
    for (int b = 0; b < (int)gridDim.x; ++b) {
        const unsigned v = __ldcg(A.hist + (size_t)b * 256 + t);
        total += v;
        if (b < (int)blockIdx.x) before += v;
    }
    unsigned incl = total;
    for (int off = 1; off < 32; off <<= 1) { const unsigned u = __shfl_up_sync(kFull, incl, off); if (lane >= off) incl += u; }
    if (lane == 31) s_scan[warp] = incl;
    __syncthreads();
    unsigned wsum = 0u;
    for (int q = 0; q < warp; ++q) wsum += s_scan[q];
    s_base[t] = wsum + incl - total + before;
    __syncthreads();
    // ---- stable scatter, 256 entries at a time in source order
    for (int i0 = first; i0 < last; i0 += kThreads) {            // CTA-uniform trip count
        const int i = i0 + t;
        const bool in = i < last;
        const unsigned k = in ? __ldcg(src_k + i) : 0u;
        const unsigned v = in ? (src_v ? __ldcg(src_v + i) : (unsigned)i) : 0u;
        const unsigned d = in ? ((k >> shift) & 255u) : (0x10000u + (unsigned)lane);      // lanes past the end: alone
#pragma unroll
        for (int q = 0; q < kWarps; ++q) s_wcount[q][t] = 0u;
        __syncthreads();
        const unsigned peers = __match_any_sync(kFull, d);
        const unsigned rank = (unsigned)__popc(peers & ((1u << lane) - 1u));
        if (in && rank == 0u) s_wcount[warp][d] = (unsigned)__popc(peers);
        __syncthreads();
        {   // digit t: the warps' counts become their starting positions, the running base moves on
            unsigned run = s_base[t];
#pragma unroll
            for (int q = 0; q < kWarps; ++q) { const unsigned c = s_wcount[q][t]; s_wcount[q][t] = run; run += c; }
            s_base[t] = run;
        }
        __syncthreads();
        if (in) {
            const unsigned pos = s_wcount[warp][d] + rank;
            dst_k[pos] = k; dst_v[pos] = v;
        }
        __syncthreads();
    }
    grid_barrier(bar_s);
}

__global__ void __launch_bounds__(kThreads) groups_kernel(Arrays A, int chunk) {
    const int n = A.n_box + A.n_sph;
    const int lane = threadIdx.x & 31;
    const int first = min(n, (int)blockIdx.x * chunk), last = min(n, first + chunk);
    // ---- 0: bounding box of the centres (NaN centres are ignored by fminf / fmaxf)
    {
        float lo[3] = {CUDART_INF_F, CUDART_INF_F, CUDART_INF_F}, hi[3] = {-CUDART_INF_F, -CUDART_INF_F, -CUDART_INF_F};
        for (int i = first + (int)threadIdx.x; i < last; i += kThreads) {
            const float4 c = bounding_record(A, i);
            lo[0] = fminf(lo[0], c.x); lo[1] = fminf(lo[1], c.y); lo[2] = fminf(lo[2], c.z);
            hi[0] = fmaxf(hi[0], c.x); hi[1] = fmaxf(hi[1], c.y); hi[2] = fmaxf(hi[2], c.z);
        }
        for (int a = 0; a < 3; ++a) {
            for (int off = 16; off > 0; off >>= 1) {
                lo[a] = fminf(lo[a], __shfl_xor_sync(kFull, lo[a], off));
                hi[a] = fmaxf(hi[a], __shfl_xor_sync(kFull, hi[a], off));
            }
            if (lane == 0 && first < last) { atomicMin(&A.sc->bb[a], enc(lo[a])); atomicMax(&A.sc->bb[3 + a], enc(hi[a])); }
        }
    }
    grid_barrier(&A.sc->sync[0]);
    // ---- 1: Morton keys, then the sort
    {
        float lo[3], scale[3];
        for (int a = 0; a < 3; ++a) {
            lo[a] = dec(__ldcg(&A.sc->bb[a]));
            const float span = dec(__ldcg(&A.sc->bb[3 + a])) - lo[a];
            scale[a] = span > 0.0f && span < CUDART_INF_F ? 255.999f / span : 0.0f;
        }
        for (int i = first + (int)threadIdx.x; i < last; i += kThreads) {
            const float4 c = bounding_record(A, i);
            const int qx = min(max((int)((c.x - lo[0]) * scale[0]), 0), 255);
            const int qy = min(max((int)((c.y - lo[1]) * scale[1]), 0), 255);
            const int qz = min(max((int)((c.z - lo[2]) * scale[2]), 0), 255);
            A.key_a[i] = spread8((unsigned)qx) | (spread8((unsigned)qy) << 1) | (spread8((unsigned)qz) << 2);
        }
    }
    __syncthreads();      // this CTA reads back only its own keys in the first pass
    radix_pass(A, A.key_a, nullptr, A.key_b, A.val_b, 0, first, last, &A.sc->sync[1], &A.sc->sync[2]);
    radix_pass(A, A.key_b, A.val_b, A.key_a, A.val_a, 8, first, last, &A.sc->sync[3], &A.sc->sync[4]);
    radix_pass(A, A.key_a, A.val_a, A.key_b, A.val_b, 16, first, last, &A.sc->sync[5], &A.sc->sync[6]);
    // ---- 2: one warp per group of 32 consecutive entries of the sorted order
    const float E = __ldcg(A.extent);
    const float pad1 = 2e-6f * E, pad2 = 4e-6f * E * E;          // as the kernels form them (rt_kernel_block.cuh)
    const int n_groups = (n + RT_GROUP_SIZE - 1) / RT_GROUP_SIZE;
    const int warps_total = (int)gridDim.x * kWarps;
    for (int g = (int)blockIdx.x * kWarps + (int)(threadIdx.x >> 5); g < n_groups; g += warps_total) {
        const int i = g * RT_GROUP_SIZE + lane;
        const bool in = i < n;
        const int hot = in ? (int)__ldcg(A.val_b + i) : 0;
        float4 c = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
        if (in) c = bounding_record(A, hot);
        // rt_cone.h: packet_radius without its last step (the packet's `grow`), the same operations in the same order
        const bool sph = hot >= A.n_box;
        float rp0 = sqrt_approx(sph ? c.w + pad2 : c.w);
        if (sph) rp0 = __fmaf_rn(rp0, 1.0005f, pad1);
        // an object whose radius is NaN (a box that can never be hit) or whose centre is not finite passes no packet test
        const bool valid = in && rp0 == rp0 && fabsf(c.x) < CUDART_INF_F && fabsf(c.y) < CUDART_INF_F && fabsf(c.z) < CUDART_INF_F;
        float lo[3] = {valid ? c.x : CUDART_INF_F, valid ? c.y : CUDART_INF_F, valid ? c.z : CUDART_INF_F};
        float hi[3] = {valid ? c.x : -CUDART_INF_F, valid ? c.y : -CUDART_INF_F, valid ? c.z : -CUDART_INF_F};
        for (int a = 0; a < 3; ++a)
            for (int off = 16; off > 0; off >>= 1) {
                lo[a] = fminf(lo[a], __shfl_xor_sync(kFull, lo[a], off));
                hi[a] = fmaxf(hi[a], __shfl_xor_sync(kFull, hi[a], off));
            }
        const bool any = __any_sync(kFull, valid);
        const float mx = 0.5f * lo[0] + 0.5f * hi[0], my = 0.5f * lo[1] + 0.5f * hi[1], mz = 0.5f * lo[2] + 0.5f * hi[2];
        const float dx = c.x - mx, dy = c.y - my, dz = c.z - mz;
        // |c - m| rounded up (1.0001) plus the member's radius; the sphere of the group holds every member's sphere
        float r = valid ? __fmaf_rn(sqrtf(dx * dx + dy * dy + dz * dz), 1.0001f, rp0) : -CUDART_INF_F;
        for (int off = 16; off > 0; off >>= 1) r = fmaxf(r, __shfl_xor_sync(kFull, r, off));
        if (lane == 0) A.grp[g] = any ? make_float4(mx, my, mz, __fmaf_rn(r, 1.0001f, 1e-5f * E)) : make_float4(0.0f, 0.0f, 0.0f, CUDART_NAN_F);
        A.srec[i] = make_float4(c.x, c.y, c.z, valid ? c.w : CUDART_NAN_F);       // NaN: passes no packet test
        A.sidx[i] = (uint32_t)hot;
    }
}

static int max_grid(int sm_count) {
    int per_sm = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, groups_kernel, kThreads, 0) != cudaSuccess || per_sm < 1) per_sm = 1;
    if (per_sm > 2) per_sm = 2;
    return per_sm * sm_count;
}
static int grid_for(int n_hot, int sm_count, int* chunk_out) {
    // every CTA resident (grid barriers); chunks of whole 256-entry rows, 1024 entries at least
    int chunk = 1024;
    const int cap = max_grid(sm_count);
    if ((n_hot + chunk - 1) / chunk > cap) chunk = ((n_hot + cap - 1) / cap + kThreads - 1) / kThreads * kThreads;
    *chunk_out = chunk;
    const int grid = (n_hot + chunk - 1) / chunk;
    return grid < 1 ? 1 : grid;
}

// bytes of scratch for n_hot objects: barrier words + bounding box, 4 key / value arrays, the histograms
size_t scratch_bytes(int n_hot, int sm_count) {
    return 256 + 4 * (((size_t)n_hot * 4 + 255) / 256 * 256) + (size_t)max_grid(sm_count) * 256 * 4;
}

// d_grp: (n_hot + 31) / 32 float4; d_srec, d_sidx: 32 entries per group.  Ordered on `st` after whatever wrote the records.
cudaError_t run(const float* d_box_cull, const float* d_sph, int n_box, int n_sph, const float* d_extent, float* d_grp, float* d_srec,
                uint32_t* d_sidx, void* d_scratch, int sm_count, cudaStream_t st) {
    const int n = n_box + n_sph;
    if (n <= 0) return cudaSuccess;
    Scratch init = {};
    for (int a = 0; a < 3; ++a) { init.bb[a] = 0xffffffffu; init.bb[3 + a] = 0u; }
    cudaError_t e = cudaMemcpyAsync(d_scratch, &init, sizeof(init), cudaMemcpyHostToDevice, st);
    if (e != cudaSuccess) return e;
    uint8_t* base = reinterpret_cast<uint8_t*>(d_scratch);
    const size_t arr = ((size_t)n * 4 + 255) / 256 * 256;
    Arrays A;
    A.box_cull = reinterpret_cast<const float4*>(d_box_cull); A.sph = reinterpret_cast<const float4*>(d_sph);
    A.n_box = n_box; A.n_sph = n_sph; A.extent = d_extent;
    A.sc = reinterpret_cast<Scratch*>(base);
    A.key_a = reinterpret_cast<unsigned int*>(base + 256);
    A.key_b = reinterpret_cast<unsigned int*>(base + 256 + arr);
    A.val_a = reinterpret_cast<unsigned int*>(base + 256 + 2 * arr);
    A.val_b = reinterpret_cast<unsigned int*>(base + 256 + 3 * arr);
    A.hist = reinterpret_cast<unsigned int*>(base + 256 + 4 * arr);
    A.grp = reinterpret_cast<float4*>(d_grp); A.srec = reinterpret_cast<float4*>(d_srec); A.sidx = d_sidx;
    int chunk = 0;
    const int grid = grid_for(n, sm_count, &chunk);
    groups_kernel<<<grid, kThreads, 0, st>>>(A, chunk);
    return cudaGetLastError();
}

}  // namespace rt_groups
