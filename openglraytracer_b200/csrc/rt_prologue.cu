/*
 * rt_prologue.cu — the host prologue as device kernels.
 *
 * The shader rebuilds local_to_world, its inverse and the normal matrix for every (ray, box) test
 * (glsl:650-652, 718); the library hoists that to once per object.  For large scenes re-uploaded every
 * frame (an animated 64k-object scene: 64k x sin/cos/inverse on one host thread + 27 MB of records
 * over PCIe) the hoisted work itself becomes the frame's bottleneck, so it moves to the GPU: the raw
 * 132-byte objects are uploaded (8.6 MB at 64k) and three small kernels build every record in HBM.
 * Same functions as the host run (rt_prologue_math.h), compiled -fmad=false; only sinf / cosf differ
 * (CUDA vs glibc), so the strict kernel keeps the host prologue.
 */
#include <cuda_runtime.h>
#include <stdint.h>

#include "rt_prologue_math.h"

namespace rt_prologue {

constexpr int kThreads = 128;

// one thread per original object: cold record + the object's hot records
__global__ void objects_kernel(const rt_object* __restrict__ objs, const int32_t* __restrict__ kind,
                               const int32_t* __restrict__ hot_index, int n, RtCold* __restrict__ cold,
                               float* __restrict__ box_fast, float* __restrict__ box_cull, float* __restrict__ sph,
                               unsigned int* __restrict__ extent_bits) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    float reach = 0.0f;
    if (i < n) {
        const rt_object o = objs[i];
        const int k = kind[i], h = hot_index[i];
        RtCold c;
        float fast12[12], cull4[4], sph4[4];
        rtpm::object_records(o, k, h, &c, fast12, nullptr, cull4, sph4);
        cold[i] = c;
        if (k == RT_KIND_BOX) {
            for (int q = 0; q < 12; q++) box_fast[(size_t)h * 12 + q] = fast12[q];
            for (int q = 0; q < 4; q++) box_cull[(size_t)h * 4 + q] = cull4[q];     // radius still un-padded
            reach = (float)rtpm::reach(cull4, fabs((double)cull4[3]));
        } else if (k == RT_KIND_SPHERE) {
            for (int q = 0; q < 4; q++) sph[(size_t)h * 4 + q] = sph4[q];
            reach = (float)rtpm::reach(sph4, sqrt(fabs((double)sph4[3])));
        }
    }
    // scene extent: max over objects (non-negative floats order like their bit patterns)
    for (int off = 16; off > 0; off >>= 1) reach = fmaxf(reach, __shfl_down_sync(0xffffffffu, reach, off));
    if ((threadIdx.x & 31) == 0 && reach > 0.0f) atomicMax(extent_bits, __float_as_uint(reach * 1.000001f));
}

// one thread per hot object (boxes then spheres): per-light records, then the padded bounding radius
__global__ void records_kernel(float* __restrict__ box_cull, const float* __restrict__ sph, int n_box, int n_sph,
                               const RtLightDev* __restrict__ lights, int n_lit, int lit0, int lit1, int lit2,
                               const unsigned int* __restrict__ extent_bits, float* __restrict__ obj_lrec) {
    const int h = blockIdx.x * blockDim.x + threadIdx.x;
    if (h >= n_box + n_sph) return;
    const double extent = (double)__uint_as_float(*extent_bits);
    const bool is_b = h < n_box;
    float c4[4];
    for (int q = 0; q < 4; q++) c4[q] = is_b ? box_cull[(size_t)h * 4 + q] : sph[(size_t)(h - n_box) * 4 + q];
    const double rad = is_b ? (double)c4[3] : sqrt((double)c4[3]);
    for (int s = 0; s < n_lit; s++) {
        const int li = s == 0 ? lit0 : (s == 1 ? lit1 : lit2);
        float rec[4];
        rtpm::light_record(c4, rad, is_b, lights[li].position, extent, rec);
        float* dst = obj_lrec + (((size_t)(h / 64) * n_lit + s) * 64 + (h % 64)) * 4;
        for (int q = 0; q < 4; q++) dst[q] = rec[q];
    }
    if (is_b) box_cull[(size_t)h * 4 + 3] = rtpm::padded_radius2((double)c4[3], extent);
}

// the padding of the last light-record tile: records that never pass
__global__ void lrec_fill_kernel(float4* __restrict__ obj_lrec, int n_f4) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n_f4) obj_lrec[i] = make_float4(0.0f, 0.0f, 0.0f, 1e30f);
}

cudaError_t run(const rt_object* d_objs, const int32_t* d_kind, const int32_t* d_hot_index, int n_obj,
                int n_box, int n_sph, RtCold* d_cold, float* d_box_fast, float* d_box_cull, float* d_sph,
                const RtLightDev* d_lights, int n_lit, const int* lit, float light_reach,
                unsigned int* d_extent_bits, float* d_obj_lrec, cudaStream_t st) {
    const unsigned int seed = __builtin_bit_cast(unsigned int, light_reach > 1.0f ? light_reach : 1.0f);
    cudaError_t e = cudaMemcpyAsync(d_extent_bits, &seed, sizeof(seed), cudaMemcpyHostToDevice, st);
    if (e != cudaSuccess) return e;
    if (n_obj > 0)
        objects_kernel<<<(n_obj + kThreads - 1) / kThreads, kThreads, 0, st>>>(d_objs, d_kind, d_hot_index, n_obj, d_cold,
                                                                               d_box_fast, d_box_cull, d_sph, d_extent_bits);
    const int n_hot = n_box + n_sph;
    if (n_lit > 0) {
        const int n_f4 = ((n_hot + 63) / 64) * n_lit * 64;
        if (n_f4 > 0) lrec_fill_kernel<<<(n_f4 + kThreads - 1) / kThreads, kThreads, 0, st>>>(reinterpret_cast<float4*>(d_obj_lrec), n_f4);
    }
    if (n_hot > 0)
        records_kernel<<<(n_hot + kThreads - 1) / kThreads, kThreads, 0, st>>>(d_box_cull, d_sph, n_box, n_sph, d_lights, n_lit,
                                                                               lit[0], lit[1], lit[2], d_extent_bits, d_obj_lrec);
    return cudaGetLastError();
}

}  // namespace rt_prologue
