/*
 * rt_kernel_block.cuh — the large-scene kernel with CTA-level packet tests (included by rt_kernel.cuh,
 * fast kernel only, inside namespace rt_fast).
 *
 * render_kernel<.., CULL> lets every warp test every object against the cone of its own work item
 * (packet_trace).  Here the 8 warps of a CTA take a BLOCK of 2 x 4 work items (48 x 16 pixels) and
 * share that pass: all 256 threads test the scene, one object per thread and trip, against the cone of
 * the whole block (primary rays: around the block's corner rays, glsl:377-392; shadow rays: per lit
 * light, around the block's hit points seen from the light, glsl:808-809) and append the objects that
 * meet it to a list in shared memory.  Each warp then runs its own cone test and the per-ray tests
 * (candidate_test) over that list only.  Below RT_GROUPS_MIN_OBJECTS every object is tested for every block
 * (the pass over the N objects is paid once per 8 work items); above, the pass asks the bounding spheres of
 * groups of 32 neighbouring objects first (rt_groups.cu, block_filter_groups) and tests only the members of
 * the groups a cone met.
 *
 * The list is unordered (warps append concurrently); candidate_test's accept rule (smaller t, then lower
 * original object index) makes the result independent of the order, so frames stay byte-identical to
 * the per-warp path and to the plain loops (tests/test_gpu_parity.py).
 *
 * Barriers: every __syncthreads() below sits in CTA-uniform control flow (conditions on kernel
 * parameters or on values read from shared memory after a barrier), never under a per-warp condition.
 */
#pragma once

constexpr int kBlkW = 2, kBlkH = 4;              // work items per block
constexpr int kListCap = 2048;                   // primary candidates per block; an overflowing block falls back to packet_trace
constexpr int kGroupCap = 1024;                  // groups tested per round of block_filter_groups
constexpr int kSplitMinList = 64;                // root shadow rays of a warp: from this many candidates in the block's list ...
constexpr float kSplitSin2 = 0.0025f;            // ... and a cone wider than sin 0.05 (2.9 degrees) the warp's hit points go in clusters
constexpr float kSplitCos = 0.99f;               // a cluster: directions from the light within 8 degrees of its leader's (0.98 .. 0.995 measure alike)
constexpr int kClusters = 12;                    // the last one takes what the others left
constexpr float kBlockMinCos = 0.1f;             // light cones of a block may be wide: they only feed the list, which cannot
                                                 // overflow (list B lives in HBM / L2, one slice of n_hot entries per CTA)

struct BlockShared {
    uint32_t list_a[kListCap];                   // primary candidates of the block: hot index | 1 << 28
    float sum[kWarps][12];                       // per warp, per light slot: sum of unit directions (xyz), count
    float mx[kWarps][12];                        // per warp, per light slot: max sin^2, max |L - p|^2, in-cone flag
    float hw[kWarps][4];                         // per warp, per light slot: max distance of a hit point to the block's fan plane
    unsigned count_a, count_b[2];                // count_b: one per sample parity (reset while the other is in use)
    unsigned count_g;                            // groups that met the block's cones (block_filter_groups)
    uint32_t list_g[kGroupCap];                  // those groups: index | cone mask << 28
    unsigned long long next;
};
constexpr int kBlockSharedBytes = (int)((sizeof(BlockShared) + 15) / 16 * 16);

__device__ __forceinline__ unsigned long long global_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
__device__ __forceinline__ unsigned smid() {
    unsigned v;
    asm volatile("mov.u32 %0, %%smid;" : "=r"(v));
    return v;
}

// All threads of the CTA: one object per thread and trip against NC cones; hits are appended to `list`.
template <bool SMEM, bool COUNT, int NC>
__device__ __forceinline__ void block_filter(const RtParams& P, const WarpCtx& w, const Cone (&cone)[NC], unsigned on,
                                             const float (&ball)[NC], float grow, float pad1, float pad2, uint32_t* __restrict__ list, int cap,
                                             unsigned* __restrict__ counter, RtCounters& cnt, float4* __restrict__ rec_list = nullptr, int rec_cap = 0) {
    const int n_hot = P.n_box + P.n_sph;
    const unsigned lt_mask = (1u << w.lane) - 1u;
    constexpr int B = 4;                          // records in flight per thread (L2 latency when streaming)
    for (int j0 = 0; j0 < n_hot; j0 += kThreads * B) {
        float4 c[B];
        int jj[B];
#pragma unroll
        for (int b = 0; b < B; ++b) {
            jj[b] = j0 + b * kThreads + (int)threadIdx.x;
            const int idx = min(jj[b], n_hot - 1);
            const bool sph = idx >= P.n_box;
            if (SMEM) c[b] = smem_f4[sph ? w.sph_off4 + (uint32_t)(idx - P.n_box) : w.box_off4 + (uint32_t)idx];
            else c[b] = sph ? __ldg(P.sph_hot + (idx - P.n_box)) : __ldg(P.box_cull + idx);
        }
#pragma unroll
        for (int b = 0; b < B; ++b) {
            if (j0 + b * kThreads >= n_hot) break;                 // CTA-uniform
            const bool in = jj[b] < n_hot;
            const bool sph = min(jj[b], n_hot - 1) >= P.n_box;
            const float rp = packet_radius(c[b].w, sph, grow, pad1, pad2);
            unsigned cm = 0u;
#pragma unroll
            for (int k = 0; k < NC; ++k) {
                if (NC == 1 || ((on >> k) & 1u)) cm |= cone_pass<NC != 1>(cone[k], c[b], rp) ? (1u << k) : 0u;
                if (ball[k] > 0.0f) cm |= ball_pass(cone[k].apex, ball[k], c[b], rp) ? (1u << k) : 0u;     // CTA-uniform branch
            }
            if (!in) cm = 0u;
            if (COUNT && in) cnt.packet_tests += (unsigned long long)(NC == 1 ? 1 : __popc(on));
            const unsigned mask = __ballot_sync(kFull, cm != 0u);
            if (mask) {
                unsigned base = 0u;
                if (w.lane == 0) base = atomicAdd(counter, (unsigned)__popc(mask));
                base = __shfl_sync(kFull, base, 0);
                const unsigned pos = base + (unsigned)__popc(mask & lt_mask);
                if (cm != 0u && pos < (unsigned)cap) list[pos] = (uint32_t)jj[b] | (cm << 28);
                if (cm != 0u && pos < (unsigned)rec_cap) rec_list[pos] = c[b];
            }
        }
    }
}

// The same list through the groups of rt_groups.cu: all threads test group spheres (one per thread and trip), then
// each warp takes a group that met a cone and tests its 32 members, one per lane, against the cones the group met --
// the member test of block_filter on a copy of the same record.  A group that no
// cone meets has no member that a cone meets (rt_cone.h: group_radius), so the list holds the same objects; their
// order differs, which candidate_test does not see.  kGroupCap groups per round; barriers in CTA-uniform flow.
template <bool COUNT, int NC>
__device__ __forceinline__ void block_filter_groups(const RtParams& P, const WarpCtx& w, BlockShared* bs, const Cone (&cone)[NC], unsigned on,
                                                    const float (&ball)[NC], float grow, float pad1, float pad2, uint32_t* __restrict__ list, int cap,
                                                    unsigned* __restrict__ counter, RtCounters& cnt, float4* __restrict__ rec_list = nullptr, int rec_cap = 0) {
    const unsigned lt_mask = (1u << w.lane) - 1u;
    const int warp = (int)(threadIdx.x >> 5);
    for (int g0 = 0; g0 < P.n_groups; g0 += kGroupCap) {
        if (threadIdx.x == 0) bs->count_g = 0u;
        __syncthreads();
        const int g1 = min(P.n_groups, g0 + kGroupCap);
#pragma unroll 1
        for (int gb = g0; gb < g1; gb += kThreads) {
            const int g = gb + (int)threadIdx.x;
            const bool in = g < g1;
            const float4 c = __ldg(P.grp + min(g, g1 - 1));
            unsigned cm = 0u;
#pragma unroll
            for (int k = 0; k < NC; ++k) {
                const bool use_c = NC == 1 || ((on >> k) & 1u), use_b = ball[k] > 0.0f;            // CTA-uniform
                if (use_c || use_b) {
                    const float rp = group_radius(cone[k].apex, c, grow);
                    if (use_c) cm |= cone_pass<NC != 1>(cone[k], c, rp) ? (1u << k) : 0u;
                    if (use_b) cm |= ball_pass(cone[k].apex, ball[k], c, rp) ? (1u << k) : 0u;
                }
            }
            if (!in) cm = 0u;
            if (COUNT && in) cnt.packet_tests += (unsigned long long)(NC == 1 ? 1 : __popc(on));
            const unsigned mask = __ballot_sync(kFull, cm != 0u);
            if (mask) {
                unsigned base = 0u;
                if (w.lane == 0) base = atomicAdd(&bs->count_g, (unsigned)__popc(mask));
                base = __shfl_sync(kFull, base, 0);
                if (cm != 0u) bs->list_g[base + (unsigned)__popc(mask & lt_mask)] = (uint32_t)g | (cm << 28);
            }
        }
        __syncthreads();
        const int ng = (int)bs->count_g;
#pragma unroll 1
        for (int e = warp; e < ng; e += kWarps) {
            const uint32_t ent = bs->list_g[e];
            const int i = (int)(ent & 0x0fffffffu) * RT_GROUP_SIZE + w.lane;
            const unsigned gm = ent >> 28;
            const float4 c = __ldg(P.srec + i);
            const uint32_t hot = __ldg(P.sidx + i);
            const float rp = packet_radius(c.w, hot >= (uint32_t)P.n_box, grow, pad1, pad2);      // NaN: no object / never hit
            unsigned cm = 0u;
#pragma unroll
            for (int k = 0; k < NC; ++k) {
                if ((gm >> k) & 1u) {                              // warp-uniform
                    if (NC == 1 || ((on >> k) & 1u)) cm |= cone_pass<NC != 1>(cone[k], c, rp) ? (1u << k) : 0u;
                    if (ball[k] > 0.0f) cm |= ball_pass(cone[k].apex, ball[k], c, rp) ? (1u << k) : 0u;
                }
            }
            if (COUNT) cnt.packet_tests += (unsigned long long)(rp == rp ? __popc(gm) : 0);
            const unsigned mask = __ballot_sync(kFull, cm != 0u);
            if (mask) {
                unsigned base = 0u;
                if (w.lane == 0) base = atomicAdd(counter, (unsigned)__popc(mask));
                base = __shfl_sync(kFull, base, 0);
                const unsigned pos = base + (unsigned)__popc(mask & lt_mask);
                if (cm != 0u && pos < (unsigned)cap) list[pos] = hot | (cm << 28);
                if (cm != 0u && pos < (unsigned)rec_cap) rec_list[pos] = c;
            }
        }
        __syncthreads();
    }
}

// One warp: its 3 rays per lane against the entries of a block list whose mask has bit `kbit`; the
// warp's own cone first (one entry per lane), then candidate_test for the survivors.
template <bool SMEM, bool COUNT, bool GLOB>
__device__ __noinline__ void list_trace(const RtParams& P, const WarpCtx& w, const uint32_t* __restrict__ list, int count, unsigned kbit,
                                        const Cone* __restrict__ cone_in, int mode, float ball_r, float grow, float live_min, float pad1, float pad2,
                                        bool shared, const V3* __restrict__ o_in, const V3* __restrict__ d_in,
                                        float* __restrict__ best_io, int* __restrict__ bidx_io, RtCounters& cnt,
                                        const float4* __restrict__ rec_list = nullptr, int rec_cap = 0,
                                        const float* __restrict__ more_cones = nullptr, int n_more = 0) {
    RaySet rs;
    rayset_load(rs, o_in, d_in, best_io, bidx_io);
    const Cone cone = cone_in[0];
    const int lane = w.lane;
    // GLOB: the list was written to global memory by this CTA during this launch: coherent (L2) loads, the next 32
    // entries in flight while the current ones are tested.  rec_list: the filter left the first rec_cap entries'
    // bounding records next to the list (one contiguous read per trip; fetching them by hot-list position is 32
    // scattered L2 reads per trip, which is what a warp's walk over a list of thousands of entries waits for).
    const int n_rec = GLOB ? min(count, rec_cap) : 0;
    uint32_t ent_next = lane < count ? (GLOB ? __ldcg(list + lane) : list[lane]) : 0u;
    float4 rec_next = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
    if (GLOB && lane < n_rec) rec_next = __ldcg(rec_list + lane);
    for (int e0 = 0; e0 < count; e0 += 32) {
        if (!__any_sync(kFull, rs.best[0] >= live_min || rs.best[1] >= live_min || rs.best[2] >= live_min)) break;
        const int e = e0 + lane;
        const uint32_t ent = ent_next;
        const float4 rec = rec_next;
        ent_next = e + 32 < count ? (GLOB ? __ldcg(list + e + 32) : list[e + 32]) : 0u;
        if (GLOB && e + 32 < n_rec) rec_next = __ldcg(rec_list + e + 32);
        const int hot = (int)(ent & 0x0fffffffu);
        const bool sel = e < count && ((ent >> 28) & kbit) != 0u;
        const bool sph = hot >= P.n_box;
        float4 c;
        if (SMEM) c = smem_f4[sph ? w.sph_off4 + (uint32_t)(hot - P.n_box) : w.box_off4 + (uint32_t)hot];
        else if (GLOB && e < n_rec) c = rec;
        else c = sph ? __ldg(P.sph_hot + (hot - P.n_box)) : __ldg(P.box_cull + hot);
        const float rp = packet_radius(c.w, sph, grow, pad1, pad2);
        // mode 1: the warp's cone and fan plane; 2: no cone (only near points);
        // ball_r > 0: plus the ball around the apex that holds the warp's near points
        bool pass = sel && ((mode == 1 && cone_pass<true>(cone, c, rp)) ||
                            (ball_r > 0.0f && ball_pass(cone.apex, ball_r, c, rp)));
        // more_cones (per-warp shared memory, 16 floats each): the warp's hit points in clusters, one cone per cluster
        for (int i = 0; i < n_more; ++i) {                        // warp-uniform; 0 for most warps
            const float* m = more_cones + 16 * i;
            Cone k2;
            k2.apex = cone.apex;
            k2.axis = mk(m[0], m[1], m[2]); k2.s = m[3]; k2.c2 = m[4]; k2.reach = m[5];
            k2.nrm = mk(m[6], m[7], m[8]); k2.hw = m[9];
            pass = pass || (sel && cone_pass<true>(k2, c, rp));
        }
        if (COUNT) cnt.packet_tests += sel;
        unsigned mask = __ballot_sync(kFull, pass);
        while (mask) {
            const int src = __ffs(mask) - 1;
            mask &= mask - 1u;
            const int hot_s = __shfl_sync(kFull, hot, src);
            const float4 cc = make_float4(__shfl_sync(kFull, c.x, src), __shfl_sync(kFull, c.y, src),
                                          __shfl_sync(kFull, c.z, src), __shfl_sync(kFull, c.w, src));
            candidate_test<COUNT>(P, rs, hot_s, cc, shared, live_min, 7u, cnt);
        }
    }
#pragma unroll
    for (int r = 0; r < kShadowBatch; ++r) { best_io[r] = rs.best[r]; bidx_io[r] = rs.bidx[r]; }
}

template <bool SMEM, bool COUNT, bool BOUNCE>
__global__ void __launch_bounds__(kThreads, RT_MIN_BLOCKS_CULL) render_block_kernel(const __grid_constant__ RtParams P) {
    constexpr bool CULL = true, REPACK = true;
    constexpr int NP = kShadowBatch;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    WarpCtx w;
    w.lane = lane;
    w.phase = 0u;
    size_t off4 = 0;                         // in float4 units; layout as in render_kernel, then BlockShared
    if (SMEM) {
        const int nb4 = P.n_box, ns4 = P.n_sph;
        for (int i = threadIdx.x; i < nb4; i += kThreads) smem_f4[i] = __ldg(P.box_cull + i);
        for (int i = threadIdx.x; i < ns4; i += kThreads) smem_f4[nb4 + i] = __ldg(P.sph_hot + i);
        w.box_off4 = 0u;
        w.sph_off4 = (uint32_t)nb4;
        off4 = (size_t)nb4 + ns4;
        w.lrec_off4 = (uint32_t)off4;
    } else {
        w.box_off4 = w.sph_off4 = w.lrec_off4 = 0u;
    }
    {
        float4* scratch = smem_f4 + off4 + (size_t)warp * (kScratchBytes / 16);
        w.queue = reinterpret_cast<float*>(scratch);
        w.occ = reinterpret_cast<int*>(scratch + kQueueRays * 2);
        off4 += (size_t)kWarps * (kScratchBytes / 16);
    }
    w.buf_off4[0] = w.buf_off4[1] = 0u;
    w.buf_addr[0] = w.buf_addr[1] = w.bar_addr[0] = w.bar_addr[1] = 0u;
    if (!SMEM) {
        w.buf_off4[0] = (uint32_t)(off4 + (size_t)warp * (2 * kTileBytes / 16));
        w.buf_off4[1] = w.buf_off4[0] + kTileBytes / 16;
        off4 += (size_t)kWarps * (2 * kTileBytes / 16);
        uint64_t* bars = reinterpret_cast<uint64_t*>(smem_f4 + off4) + warp * 2;
        off4 += (size_t)kWarps;               // 16 bytes of barriers per warp
        w.buf_addr[0] = smem_addr(smem_f4 + w.buf_off4[0]); w.buf_addr[1] = smem_addr(smem_f4 + w.buf_off4[1]);
        w.bar_addr[0] = smem_addr(bars); w.bar_addr[1] = smem_addr(bars + 1);
        if (lane == 0) {
            asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(w.bar_addr[0]) : "memory");
            asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(w.bar_addr[1]) : "memory");
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        }
    }
    BlockShared* bs = reinterpret_cast<BlockShared*>(smem_f4 + off4);
    uint32_t* __restrict__ list_b = P.blist + (size_t)blockIdx.x * P.blist_stride;    // shadow candidates: hot index | lit-light mask << 28
    const int rec_cap = SMEM ? 0 : P.brec_cap;                                         // ... and the bounding records of the first rec_cap
    float4* __restrict__ rec_b = rec_cap > 0 ? P.brec + (size_t)blockIdx.x * P.brec_cap : nullptr;
    if (threadIdx.x == 0) { bs->count_a = 0u; bs->count_b[0] = 0u; bs->count_b[1] = 0u; }
    __syncthreads();

    const int lx = lane & 7, ly = lane >> 3;
    const float E = __ldg(P.extent);
    const float pk_pad1 = 2e-6f * E, pk_pad2 = 4e-6f * E * E;
    const int bx_n = (P.tiles_x + kBlkW - 1) / kBlkW, by_n = (P.tiles_y + kBlkH - 1) / kBlkH;
    const long long blocks_per_frame = (long long)bx_n * by_n;
    const long long n_blocks = blocks_per_frame * P.n_frames;
    const int n_lights_b = (P.n_lit_batch > 0 && P.n_lit_batch <= kShadowBatch) ? P.n_lit_batch : 0;    // CTA-uniform

    RtCounters cnt;
    cnt.packet_tests = cnt.group_passes = cnt.loop_passes = cnt.candidates = cnt.shadow_group_passes = 0ull;
    cnt.rays = cnt.nodes = cnt.hit_nodes = cnt.shadow_rays = cnt.box_tests = cnt.sphere_tests = cnt.red_pixels = cnt.pad = 0ull;
    cnt.cull_tests = cnt.warp_iters = cnt.shaded_lights = 0ull;

    for (long long blk = blockIdx.x; blk < n_blocks;) {
        // RT_BLOCK_LOG (instrumented twin only): when this block started, how long each warp worked, list sizes
        const bool logging = COUNT && P.blk_log != nullptr;
        unsigned long long log_t0 = 0ull;
        unsigned log_nb = 0u, log_flags = 0u;
        if (logging) log_t0 = global_ns();
        const int frame = (int)(blk / blocks_per_frame);
        const int bin = (int)(blk - (long long)frame * blocks_per_frame);
        const int by = bin / bx_n, bx = bin - by * bx_n;
        const int tx = bx * kBlkW + (warp & (kBlkW - 1)), ty = by * kBlkH + (warp / kBlkW);     // this warp's work item
        const bool item_ok = tx < P.tiles_x && ty < P.tiles_y;
        const int py = P.y0 + ty * 4 + ly;
        int px[NP]; bool valid[NP];
#pragma unroll
        for (int q = 0; q < NP; ++q) {
            px[q] = (tx * NP + q) * 8 + lx;
            valid[q] = item_ok && px[q] < P.width && py < P.y1;
        }
        const RtCameraDev* __restrict__ cam = P.cams ? P.cams + frame : &P.cam0;
        const V3 org = mk(cam->position[0], cam->position[1], cam->position[2]);   // glsl:391

        // ---- the block's primary cone: axis through the middle of its 48 x 16 pixels, half-angle from its corner
        // rays (the angle to the axis is quasi-convex on the image plane, so a corner attains the maximum)
        Cone bc[1];
        bool a_ok;
        {
            const int x0 = bx * kBlkW * 8 * NP, yb = P.y0 + by * kBlkH * 4;
            bc[0].apex = org;
            bc[0].axis = primary_dir(P, cam, x0 + kBlkW * 4 * NP, yb + kBlkH * 2, 0.0f, 0.0f);
            const V3 cd = primary_dir(P, cam, x0 + (lane & 1) * kBlkW * 8 * NP, yb + ((lane >> 1) & 1) * kBlkH * 4, 0.0f, 0.0f);
            const V3 x = cross3(bc[0].axis, cd);
            const float s2 = fminf(__fmaf_rn(x.x, x.x, __fmaf_rn(x.y, x.y, __fmul_rn(x.z, x.z))), 1.0f);
            a_ok = __all_sync(kFull, dot3(bc[0].axis, cd) >= kConeMinCos);       // same inputs in every warp: CTA-uniform
            cone_close(bc[0], warp_max_nonneg(s2), kFar * 1.001f);
        }
        const float no_ball[1] = {0.0f};
        if (a_ok) {                                                             // CTA-uniform, as is P.n_groups
            if (P.n_groups > 0) block_filter_groups<COUNT, 1>(P, w, bs, bc, 1u, no_ball, 0.0f, pk_pad1, pk_pad2, bs->list_a, kListCap, &bs->count_a, cnt);
            else block_filter<SMEM, COUNT, 1>(P, w, bc, 1u, no_ball, 0.0f, pk_pad1, pk_pad2, bs->list_a, kListCap, &bs->count_a, cnt);
        }
        __syncthreads();                                                        // list A complete
        const int na = (int)bs->count_a;
        const bool use_a = a_ok && na <= kListCap;
        const unsigned long long log_ta = logging ? global_ns() : 0ull;

        V3 sum[NP];
#pragma unroll
        for (int q = 0; q < NP; ++q) sum[q] = mk(0.0f, 0.0f, 0.0f);
        for (int s = 0; s < P.spp; ++s) {
            V3 dir[kShadowBatch];
#pragma unroll
            for (int q = 0; q < NP; ++q) {
                float jx, jy;
                jitter(P, px[q], py, s, jx, jy);
                dir[q] = primary_dir(P, cam, px[q], py, jx, jy);
            }
            // PHASE 0 of the three root frames (glsl:956-967): the warp's own cone, then the block list
            V3 o3[kShadowBatch];
            float best[kShadowBatch]; int bidx[kShadowBatch];
#pragma unroll
            for (int q = 0; q < kShadowBatch; ++q) { o3[q] = org; bidx[q] = -1; best[q] = valid[q] ? kFar : 0.0f; }
            Cone pc[1];
            bool packet;
            {
                pc[0].apex = org;
                pc[0].axis = primary_dir(P, cam, tx * (8 * NP) + 4 * NP, P.y0 + ty * 4 + 2, 0.0f, 0.0f);
                float s2 = 0.0f;
                bool in_cone = true;
#pragma unroll
                for (int q = 0; q < kShadowBatch; ++q) {
                    const V3 x = cross3(pc[0].axis, dir[q]);
                    const float s2q = __fmaf_rn(x.x, x.x, __fmaf_rn(x.y, x.y, __fmul_rn(x.z, x.z)));
                    if (valid[q]) {
                        s2 = fmaxf(s2, fminf(s2q, 1.0f));
                        in_cone = in_cone && dot3(pc[0].axis, dir[q]) >= kConeMinCos;
                    }
                }
                packet = __all_sync(kFull, in_cone);
                cone_close(pc[0], warp_max_nonneg(s2), kFar * 1.001f);
            }
            if (packet && use_a) {
                list_trace<SMEM, COUNT, false>(P, w, bs->list_a, na, 1u, pc, 1, 0.0f, 0.0f, 1e-45f, pk_pad1, pk_pad2, true, o3, dir, best, bidx, cnt);
            } else if (packet) {
                packet_trace<SMEM, COUNT, 1>(P, w, pc, 1u, 0.0f, 1e-45f, pk_pad1, pk_pad2, true, o3, dir, best, bidx, cnt);
            } else {
                const int n_obj_tiles = n_box_tiles<CULL>(P) + n_sph_tiles(P);
                tile_begin<SMEM, CULL>(P, w);
                for (int t = 0; t < n_obj_tiles; ++t) {
                    const float4* recs = tile_acquire<SMEM, CULL>(P, w, t, n_obj_tiles);
                    any_tile<kShadowBatch, false, true, COUNT, CULL>(P, recs, t, o3, dir, best, bidx, cnt);
                }
            }

            // ---- root hits: hit point, shadow-ray origin and "needs shading" exactly as trace_sample builds them
            bool have_sb = false;
            float rsb[kShadowBatch][kShadowBatch];       // [tile q][light slot]
            V3 hp[kShadowBatch], hn3[kShadowBatch];      // hit point and shading normal of the three root hits
            int ob3[kShadowBatch];                       // their objects' original indices (-1: no hit)
#pragma unroll
            for (int q = 0; q < kShadowBatch; ++q) { hp[q] = mk(0.0f, 0.0f, 0.0f); hn3[q] = mk(0.0f, 0.0f, 1.0f); ob3[q] = -1; }
#pragma unroll
            for (int q = 0; q < kShadowBatch; ++q)
#pragma unroll
                for (int k = 0; k < kShadowBatch; ++k) rsb[q][k] = 0.0f;
            if (n_lights_b > 0) {                        // CTA-uniform
                V3 so3[kShadowBatch];
                bool sh[kShadowBatch];
#pragma unroll 1
                for (int q = 0; q < kShadowBatch; ++q) {
                    const V3 dq = q == 0 ? dir[0] : (q == 1 ? dir[1] : dir[2]);
                    const float tq = q == 0 ? best[0] : (q == 1 ? best[1] : best[2]);
                    const int iq = q == 0 ? bidx[0] : (q == 1 ? bidx[1] : bidx[2]);
                    const bool vq = q == 0 ? valid[0] : (q == 1 ? valid[1] : valid[2]);
                    bool shade = false;
                    V3 p = mk(0.0f, 0.0f, 0.0f), so = mk(0.0f, 0.0f, 0.0f), nn = mk(0.0f, 0.0f, 1.0f);
                    int ob = -1;
                    if (vq && iq >= 0) {
                        const Hit h = resolve_hit<SMEM, CULL>(P, w, org, dq, tq, iq);
                        const RtCold* __restrict__ mat = P.cold + h.obj;
                        float wp = 1.0f;                 // the weight of the Phong term at the root node
                        if (P.max_depth > 0 && mat->reflectivity > 0.0f) wp = wp * (1.0f - mat->reflectivity);
                        if (P.max_depth > 0 && mat->transparency > 0.0f) wp = wp * (1.0f - mat->transparency);
                        shade = wp != 0.0f;
                        p = h.p; nn = h.n; ob = h.obj;
                        so = axpy3(h.n, 0.01f, h.p);                                            // glsl:808
                    }
                    if (q == 0) { hp[0] = p; hn3[0] = nn; ob3[0] = ob; so3[0] = so; sh[0] = shade; }
                    else if (q == 1) { hp[1] = p; hn3[1] = nn; ob3[1] = ob; so3[1] = so; sh[1] = shade; }
                    else { hp[2] = p; hn3[2] = nn; ob3[2] = ob; so3[2] = so; sh[2] = shade; }
                }
                // ---- the block's light cones: sum of the unit directions light -> hit point over the CTA ...
                V3 lp[kShadowBatch];
#pragma unroll
                for (int k = 0; k < kShadowBatch; ++k) {
                    const RtLightDev* __restrict__ L = P.lights + P.lit[k < n_lights_b ? k : 0];
                    lp[k] = mk(L->position[0], L->position[1], L->position[2]);
                }
                // Hit points much closer to the light than the rest (an object next to a lamp, seen in front of a far
                // wall) would blow the cone up; they go into a ball around the light instead: a shadow ray that
                // starts within r of L and runs toward L stays within r of L.  near = closer than a quarter of the
                // warp's mean distance; each warp may classify as it likes, every point lands in the cone or the ball.
                // (unit directions and distances are recomputed where they are needed instead of living in 45 registers
                // across the CTA barriers; `thr` is all a later pass needs to classify a point again)
                float thr[kShadowBatch];
                auto light_vec = [&](int k, int q, V3& uq, float& l2q) {
                    const V3 v = hp[q] - lp[k];
                    l2q = __fmaf_rn(v.x, v.x, __fmaf_rn(v.y, v.y, __fmul_rn(v.z, v.z)));
                    uq = v * rsqrtf(l2q);
                };
#pragma unroll
                for (int k = 0; k < kShadowBatch; ++k) {
                    thr[k] = 0.0f;
                    if (k >= n_lights_b) continue;                     // CTA-uniform: no such lit light (the built-in lights: two)
                    float dsum = 0.0f, dn = 0.0f;
                    V3 u[kShadowBatch]; float l2[kShadowBatch];
#pragma unroll
                    for (int q = 0; q < kShadowBatch; ++q) {
                        light_vec(k, q, u[q], l2[q]);
                        if (sh[q]) { dsum += sqrt_approx(l2[q]); dn += 1.0f; }
                    }
#pragma unroll
                    for (int off = 16; off > 0; off >>= 1) {
                        dsum += __shfl_xor_sync(kFull, dsum, off);
                        dn += __shfl_xor_sync(kFull, dn, off);
                    }
                    thr[k] = dn > 0.0f ? 0.25f * dsum / dn : 0.0f;
                    float acc[4] = {0.0f, 0.0f, 0.0f, 0.0f};
#pragma unroll
                    for (int q = 0; q < kShadowBatch; ++q) {
                        const bool nearq = sh[q] && l2[q] < thr[k] * thr[k];
                        if (sh[q] && !nearq && k < n_lights_b) { acc[0] += u[q].x; acc[1] += u[q].y; acc[2] += u[q].z; acc[3] += 1.0f; }
                    }
#pragma unroll
                    for (int off = 16; off > 0; off >>= 1) {
#pragma unroll
                        for (int c4 = 0; c4 < 4; ++c4) acc[c4] += __shfl_xor_sync(kFull, acc[c4], off);
                    }
                    if (lane == 0) {
#pragma unroll
                        for (int c4 = 0; c4 < 4; ++c4) bs->sum[warp][4 * k + c4] = acc[c4];
                    }
                }
                __syncthreads();                                                    // sums visible
                Cone lc[kShadowBatch];
                unsigned on = 0u;
#pragma unroll
                for (int k = 0; k < kShadowBatch; ++k) {
                    lc[k].apex = lp[k]; lc[k].axis = mk(0.0f, 0.0f, 1.0f);
                    if (k >= n_lights_b) continue;                     // unused slot: its `on` bit stays off
                    float t4[4] = {0.0f, 0.0f, 0.0f, 0.0f};
                    for (int ww = 0; ww < kWarps; ++ww) {
#pragma unroll
                        for (int c4 = 0; c4 < 4; ++c4) t4[c4] += bs->sum[ww][4 * k + c4];
                    }
                    lc[k].apex = lp[k];
                    lc[k].axis = normalize3(mk(t4[0], t4[1], t4[2]));
                    if (k < n_lights_b && t4[3] > 0.0f) on |= 1u << k;
                    // ... then the half-angle, the reach and "all inside kBlockMinCos" over the CTA (far points), and the
                    // radius of the ball (near points)
                    float s2 = 0.0f, lmax = 0.0f, nmax = 0.0f, hmax = 0.0f;
                    bool in_cone = true;
                    const V3 bn = fan_normal(bc[0].axis, org - lp[k]);      // the block's fan plane: through the light and the block's view axis
#pragma unroll
                    for (int q = 0; q < kShadowBatch; ++q) {
                        V3 uq; float l2q;
                        light_vec(k, q, uq, l2q);
                        const bool nearq = sh[q] && l2q < thr[k] * thr[k];
                        const V3 x = cross3(lc[k].axis, uq);
                        const float s2q = __fmaf_rn(x.x, x.x, __fmaf_rn(x.y, x.y, __fmul_rn(x.z, x.z)));
                        if (sh[q] && !nearq) {
                            s2 = fmaxf(s2, fminf(s2q, 1.0f));
                            lmax = fmaxf(lmax, l2q);
                            in_cone = in_cone && dot3(lc[k].axis, uq) >= kBlockMinCos;    // NaN -> false
                            hmax = fmaxf(hmax, fabsf(dot3(bn, hp[q] - lp[k])));
                        }
                        if (nearq && k < n_lights_b) nmax = fmaxf(nmax, l2q);
                    }
                    const float s2w = warp_max_nonneg(s2), lw = warp_max_nonneg(lmax), nw = warp_max_nonneg(nmax);
                    const float hwp = warp_max_nonneg(hmax < 1e30f ? hmax : 1e30f);                // NaN -> no plane
                    const bool okw = __all_sync(kFull, in_cone);
                    if (lane == 0) { bs->mx[warp][4 * k] = s2w; bs->mx[warp][4 * k + 1] = lw; bs->mx[warp][4 * k + 2] = okw ? 1.0f : 0.0f; bs->mx[warp][4 * k + 3] = nw; bs->hw[warp][k] = hwp; }
                }
                __syncthreads();                                                    // maxima visible
                float ball[kShadowBatch];
#pragma unroll
                for (int k = 0; k < kShadowBatch; ++k) {
                    if (k >= n_lights_b) { cone_close(lc[k], 0.0f, 0.0f); ball[k] = 0.0f; continue; }
                    float s2 = 0.0f, lmax = 0.0f, nmax = 0.0f, hmax = 0.0f;
                    bool k_ok = true;
                    for (int ww = 0; ww < kWarps; ++ww) {
                        s2 = fmaxf(s2, bs->mx[ww][4 * k]);
                        lmax = fmaxf(lmax, bs->mx[ww][4 * k + 1]);
                        nmax = fmaxf(nmax, bs->mx[ww][4 * k + 3]);
                        hmax = fmaxf(hmax, bs->hw[ww][k]);
                        k_ok = k_ok && bs->mx[ww][4 * k + 2] != 0.0f;
                    }
                    cone_close(lc[k], s2, __fmaf_rn(sqrt_approx(lmax), 1.0001f, 0.0101f));
                    // hit points spread beyond kBlockMinCos as seen from the light: no angular bound, the reach and the
                    // fan plane still hold
                    if (!k_ok) cone_open(lc[k]);
                    cone_fan(lc[k], fan_normal(bc[0].axis, org - lp[k]), P.fan ? hmax : 3e38f, pk_pad1);
                    ball[k] = nmax > 0.0f ? __fmaf_rn(sqrt_approx(nmax), 1.0001f, 0.0101f) : 0.0f;
                }
                const bool any_b = on != 0u || ball[0] > 0.0f || ball[1] > 0.0f || ball[2] > 0.0f;
                // every thread holds the same `on`, balls and cones (all read from shared memory): CTA-uniform.  The list
                // cannot overflow: one entry per object at most, and the CTA's slice holds n_hot entries.
                if (any_b) {
                    if (P.n_groups > 0) block_filter_groups<COUNT, kShadowBatch>(P, w, bs, lc, on, ball, 0.0101f, pk_pad1, pk_pad2, list_b, P.blist_stride, &bs->count_b[s & 1], cnt, rec_b, rec_cap);
                    else block_filter<SMEM, COUNT, kShadowBatch>(P, w, lc, on, ball, 0.0101f, pk_pad1, pk_pad2, list_b, P.blist_stride, &bs->count_b[s & 1], cnt, rec_b, rec_cap);
                }
                __syncthreads();                                                    // list B complete
                const int nb = (int)bs->count_b[s & 1];
                if (threadIdx.x == 0) bs->count_b[(s + 1) & 1] = 0u;                // its readers (sample s - 1) are past two barriers
                if (logging) { log_nb = (unsigned)nb; log_flags |= 1u | (on << 4) | (ball[0] > 0.0f ? 0x100u : 0u) | (ball[1] > 0.0f ? 0x200u : 0u) | (ball[2] > 0.0f ? 0x400u : 0u); }

                // ---- this warp's root shadow rays, light by light (one cone over its own 96 hit points)
                if (packet) {
                    have_sb = true;
                    if (__any_sync(kFull, sh[0] || sh[1] || sh[2])) {
#pragma unroll 1
                        for (int k = 0; k < n_lights_b; ++k) {
                            const V3 lpk = k == 0 ? lp[0] : (k == 1 ? lp[1] : lp[2]);
                            V3 sd3[kShadowBatch], uu[kShadowBatch];
                            float ll[kShadowBatch], sbk[kShadowBatch];
                            bool nq[kShadowBatch];
                            V3 sm = mk(0.0f, 0.0f, 0.0f);
#pragma unroll
                            for (int q = 0; q < kShadowBatch; ++q) {
                                sd3[q] = sh[q] ? mk(lpk.x - hp[q].x, lpk.y - hp[q].y, lpk.z - hp[q].z) : mk(0.0f, 0.0f, 1.0f);   // glsl:809
                                sbk[q] = sh[q] ? 1.0f : 0.0f;
                                ll[q] = __fmaf_rn(sd3[q].x, sd3[q].x, __fmaf_rn(sd3[q].y, sd3[q].y, __fmul_rn(sd3[q].z, sd3[q].z)));
                                uu[q] = sd3[q] * -rsqrtf(ll[q]);
                                const float thk = k == 0 ? thr[0] : (k == 1 ? thr[1] : thr[2]);
                                nq[q] = sh[q] && ll[q] < thk * thk;
                                if (sh[q] && !nq[q]) sm = sm + uu[q];
                            }
#pragma unroll
                            for (int off = 16; off > 0; off >>= 1) {
                                sm.x += __shfl_xor_sync(kFull, sm.x, off);
                                sm.y += __shfl_xor_sync(kFull, sm.y, off);
                                sm.z += __shfl_xor_sync(kFull, sm.z, off);
                            }
                            Cone wc[1];
                            wc[0].apex = lpk;
                            wc[0].axis = normalize3(sm);
                            float s2 = 0.0f, lmax = 0.0f, nmax = 0.0f, hmax = 0.0f;
                            bool in_cone = true, far_pt = false;
                            const V3 wn = fan_normal(pc[0].axis, org - lpk);     // the work item's fan plane
#pragma unroll
                            for (int q = 0; q < kShadowBatch; ++q) {
                                const V3 x = cross3(wc[0].axis, uu[q]);
                                const float s2q = __fmaf_rn(x.x, x.x, __fmaf_rn(x.y, x.y, __fmul_rn(x.z, x.z)));
                                if (sh[q] && !nq[q]) {
                                    far_pt = true;
                                    s2 = fmaxf(s2, fminf(s2q, 1.0f));
                                    lmax = fmaxf(lmax, ll[q]);
                                    in_cone = in_cone && dot3(wc[0].axis, uu[q]) >= kConeMinCos;
                                    hmax = fmaxf(hmax, fabsf(dot3(wn, sd3[q])));
                                }
                                if (nq[q]) nmax = fmaxf(nmax, ll[q]);
                            }
                            const bool narrow = __all_sync(kFull, in_cone);
                            const bool has_far = __any_sync(kFull, far_pt);
                            const float nw = warp_max_nonneg(nmax);
                            const float ball_r = nw > 0.0f ? __fmaf_rn(sqrt_approx(nw), 1.0001f, 0.0101f) : 0.0f;
                            const float s2w = warp_max_nonneg(s2);
                            int n_more = 0;
                            // A warp whose 96 hit points lie at several depths (it looks past a near object at far ones) sees
                            // them under a wide angle from the light: one cone around all of them meets hundreds of the
                            // block's candidates, and each costs a per-ray test (blocks holding such warps took 600 us at 64k
                            // objects against a median of 85 and set the kernel's tail).  Such a warp puts its hit points in up
                            // to kClusters clusters -- a leader's direction from the light and everything within kSplitCos of
                            // it, the last cluster takes what is left -- and a list entry must meet one of the clusters' cones
                            // (one walk over the list; 3 clusters: 16k objects 0.765 -> 0.727 ms, up to 12: 0.63 ms).
                            if (nb >= kSplitMinList && has_far && P.split_cones && (!narrow || s2w > kSplitSin2)) {
                                bool todo[kShadowBatch];
#pragma unroll
                                for (int q = 0; q < kShadowBatch; ++q) todo[q] = sh[q] && !nq[q];
                                int n_cl = 0;
                                // every cone costs each trip of the walk a test: one cluster per 32 list entries, 2..kClusters
                                const int max_cl = min(kClusters, max(2, nb >> 5));
#pragma unroll 1
                                for (int round = 0; round < max_cl; ++round) {
                                    const unsigned r0 = __ballot_sync(kFull, todo[0]), r1 = __ballot_sync(kFull, todo[1]), r2 = __ballot_sync(kFull, todo[2]);
                                    if ((r0 | r1 | r2) == 0u) break;
                                    const int ql = r0 ? 0 : (r1 ? 1 : 2);
                                    const int lead = __ffs(r0 ? r0 : (r1 ? r1 : r2)) - 1;
                                    const V3 us = ql == 0 ? uu[0] : (ql == 1 ? uu[1] : uu[2]);
                                    const V3 lu = mk(__shfl_sync(kFull, us.x, lead), __shfl_sync(kFull, us.y, lead), __shfl_sync(kFull, us.z, lead));
                                    const bool last = round == max_cl - 1;
                                    bool in_g[kShadowBatch];
                                    V3 gs = mk(0.0f, 0.0f, 0.0f);
#pragma unroll
                                    for (int q = 0; q < kShadowBatch; ++q) {
                                        in_g[q] = todo[q] && (last || dot3(uu[q], lu) >= kSplitCos);
                                        if (in_g[q]) gs = gs + uu[q];
                                        todo[q] = todo[q] && !in_g[q];
                                    }
#pragma unroll
                                    for (int off = 16; off > 0; off >>= 1) {
                                        gs.x += __shfl_xor_sync(kFull, gs.x, off);
                                        gs.y += __shfl_xor_sync(kFull, gs.y, off);
                                        gs.z += __shfl_xor_sync(kFull, gs.z, off);
                                    }
                                    Cone gc;
                                    gc.apex = lpk;
                                    gc.axis = normalize3(gs);
                                    float g_s2 = 0.0f, g_l = 0.0f, g_h = 0.0f;
                                    bool g_in = true;
#pragma unroll
                                    for (int q = 0; q < kShadowBatch; ++q) {
                                        const V3 x = cross3(gc.axis, uu[q]);
                                        if (in_g[q]) {
                                            g_s2 = fmaxf(g_s2, fminf(__fmaf_rn(x.x, x.x, __fmaf_rn(x.y, x.y, __fmul_rn(x.z, x.z))), 1.0f));
                                            g_l = fmaxf(g_l, ll[q]);
                                            g_in = g_in && dot3(gc.axis, uu[q]) >= kConeMinCos;
                                            g_h = fmaxf(g_h, fabsf(dot3(wn, sd3[q])));
                                        }
                                    }
                                    cone_close(gc, warp_max_nonneg(g_s2), __fmaf_rn(sqrt_approx(warp_max_nonneg(g_l)), 1.0001f, 0.0101f));
                                    if (!__all_sync(kFull, g_in)) cone_open(gc);
                                    cone_fan(gc, wn, P.fan ? warp_max_nonneg(g_h < 1e30f ? g_h : 1e30f) : 3e38f, pk_pad1);
                                    if (n_cl == 0) {
                                        wc[0] = gc;
                                    } else if (lane == 0) {
                                        float* m = w.queue + 16 * (n_cl - 1);
                                        m[0] = gc.axis.x; m[1] = gc.axis.y; m[2] = gc.axis.z; m[3] = gc.s; m[4] = gc.c2; m[5] = gc.reach;
                                        m[6] = gc.nrm.x; m[7] = gc.nrm.y; m[8] = gc.nrm.z; m[9] = gc.hw;
                                    }
                                    ++n_cl;
                                }
                                __syncwarp();
                                n_more = n_cl - 1;
                            } else {
                            cone_close(wc[0], s2w, __fmaf_rn(sqrt_approx(warp_max_nonneg(lmax)), 1.0001f, 0.0101f));
                            if (!narrow) cone_open(wc[0]);                                // reach and fan plane only
                            cone_fan(wc[0], wn, P.fan ? warp_max_nonneg(hmax < 1e30f ? hmax : 1e30f) : 3e38f, pk_pad1);
                            }
                            int unused[kShadowBatch] = {-1, -1, -1};
                            list_trace<SMEM, COUNT, true>(P, w, list_b, nb, 1u << k, wc, has_far ? 1 : 2, ball_r, 0.0101f, 1.0f, pk_pad1, pk_pad2, false, so3, sd3, sbk, unused, cnt, rec_b, rec_cap, w.queue, n_more);
                            __syncwarp();
#pragma unroll
                            for (int q = 0; q < kShadowBatch; ++q) {
                                if (k == 0) rsb[q][0] = sbk[q]; else if (k == 1) rsb[q][1] = sbk[q]; else rsb[q][2] = sbk[q];
                            }
                        }
                    }
                }
            }
            // bounce rays: the three trees of a lane in lockstep, rays grouped over all 96 (trace_sample3)
            bool lockstep = false;
            if constexpr (BOUNCE) {
                lockstep = P.group_k > 0.0f && P.n_lit_batch >= 0;                        // CTA-uniform
                if (lockstep) {
                    V3 c3[kShadowBatch];
                    trace_sample3<SMEM, COUNT>(P, w, valid, org, dir, best, bidx, have_sb, rsb, pk_pad1, pk_pad2, cnt, c3);
#pragma unroll
                    for (int q = 0; q < NP; ++q) sum[q] = sum[q] + c3[q];
                }
            }
            // MAX_RAYTRACE_DEPTH = 0 with the root shadow rays already traced above: the tree is its root, whose hit is
            // resolved and whose lights are known -- ads_phong_lighting (glsl:789-840) right here, through the helpers
            // trace_sample uses (phong_light / phong_accumulate, weight 1), so the frame is the same byte for byte.
            bool shaded_here = false;
            if constexpr (!BOUNCE) {
                shaded_here = have_sb;                                                    // CTA-uniform
                if (shaded_here) {
#pragma unroll
                    for (int q = 0; q < NP; ++q) {
                        const bool hit = valid[q] && ob3[q] >= 0;
                        if (COUNT && valid[q]) { cnt.rays++; cnt.nodes++; cnt.hit_nodes += hit; }
                        if (hit) {
                            const RtCold* __restrict__ mat = P.cold + ob3[q];
                            if (COUNT) { cnt.rays += P.n_lights; cnt.shadow_rays += P.n_lights; cnt.pad += P.n_lights - P.n_lit_batch; }
                            float amb[4], dif[4] = {0, 0, 0, 0}, spc[4] = {0, 0, 0, 0};
#pragma unroll
                            for (int k = 0; k < 4; ++k) amb[k] = P.amb_sum[k] * mat->ambient[k];      // glsl:801
                            const V3 view = normalize3(-dir[q]);                                   // glsl:827
#pragma unroll
                            for (int k = 0; k < kShadowBatch; ++k) {
                                if (k < n_lights_b && !(rsb[q][k] < 1.0f)) {                       // glsl:816
                                    if (COUNT) cnt.shaded_lights++;
                                    const RtLightDev* __restrict__ L = P.lights + P.lit[k];
                                    phong_light(L, mat, mk(L->position[0] - hp[q].x, L->position[1] - hp[q].y, L->position[2] - hp[q].z),
                                                hn3[q], view, dif, spc);
                                }
                            }
                            sum[q] = sum[q] + phong_accumulate(mk(0.0f, 0.0f, 0.0f), 1.0f, amb, dif, spc, mat);
                        }
                    }
                    if (COUNT) cnt.warp_iters += NP;
                }
            }
            // one copy of the bounce machinery in the instruction stream: rolled loop, selected operands
#pragma unroll 1
            for (int q = 0; q < ((lockstep || shaded_here) ? 0 : NP); ++q) {
                const V3 dq = q == 0 ? dir[0] : (q == 1 ? dir[1] : dir[2]);
                const float tq = q == 0 ? best[0] : (q == 1 ? best[1] : best[2]);
                const int iq = q == 0 ? bidx[0] : (q == 1 ? bidx[1] : bidx[2]);
                const bool vq = q == 0 ? valid[0] : (q == 1 ? valid[1] : valid[2]);
                float fsb[kShadowBatch];
#pragma unroll
                for (int k = 0; k < kShadowBatch; ++k) fsb[k] = q == 0 ? rsb[0][k] : (q == 1 ? rsb[1][k] : rsb[2][k]);
                const V3 c = trace_sample<SMEM, REPACK, COUNT, CULL, BOUNCE>(P, w, vq, org, dq, true, tq, iq, have_sb, fsb, pk_pad1, pk_pad2, cnt);
                if (q == 0) sum[0] = sum[0] + c;
                else if (q == 1) sum[1] = sum[1] + c;
                else sum[2] = sum[2] + c;
            }
        }
        uint8_t* frame_out = P.out + (long long)frame * P.frame_stride;
#pragma unroll
        for (int q = 0; q < NP; ++q) {
            const float fs = (float)P.spp;
            const V3 fc = mk(fdiv(sum[q].x, fs), fdiv(sum[q].y, fs), fdiv(sum[q].z, fs));
            const uint32_t word = unorm8(fc.x) | (unorm8(fc.y) << 8) | (unorm8(fc.z) << 16);      // glsl:404
            const long long pix = (long long)(py - P.out_row0) * P.width + px[q];
            const uint32_t w1 = __shfl_down_sync(kFull, word, 1);
            const uint32_t w2 = __shfl_down_sync(kFull, word, 2);
            const uint32_t w3 = __shfl_down_sync(kFull, word, 3);
            const bool vec_ok = ((P.width & 3) == 0) && ((tx * NP + q) * 8 + 8 <= P.width) &&
                                ((reinterpret_cast<uintptr_t>(frame_out) & 15) == 0);
            if (vec_ok) {
                if ((lane & 3) == 0 && item_ok && py < P.y1)
                    *reinterpret_cast<uint4*>(frame_out + pix * 4) = make_uint4(word, w1, w2, w3);
            } else if (valid[q]) {
                *reinterpret_cast<uint32_t*>(frame_out + pix * 4) = word;
            }
        }
        if (logging) {
            uint32_t* rec = P.blk_log + blk * 16;
            const unsigned long long t1 = global_ns();
            if (lane == 0) rec[8 + warp] = (uint32_t)(t1 - log_t0);
            if (threadIdx.x == 0) {
                rec[0] = (uint32_t)log_t0; rec[1] = (uint32_t)(log_t0 >> 32);
                rec[2] = (uint32_t)(log_ta - log_t0); rec[3] = (uint32_t)na; rec[4] = log_nb;
                rec[5] = log_flags | (use_a ? 4u : 8u); rec[6] = smid(); rec[7] = blockIdx.x;
            }
        }
        // next block for the whole CTA; list A is reset for it (every warp is past its reads of both lists here
        // only after the barrier, so the reset sits behind it)
        if (threadIdx.x == 0) bs->next = (unsigned long long)gridDim.x + atomicAdd(P.tile_counter, 1ull);
        __syncthreads();
        blk = (long long)bs->next;
        if (threadIdx.x == 0) { bs->count_a = 0u; bs->count_b[0] = 0u; bs->count_b[1] = 0u; }
        __syncthreads();
    }

    if (COUNT) {
        if (lane != 0) cnt.warp_iters = 0ull;    // a per-warp count, kept by lane 0
        unsigned long long v[16] = {cnt.rays, cnt.nodes, cnt.hit_nodes, cnt.shadow_rays, cnt.box_tests, cnt.sphere_tests,
                                    cnt.red_pixels, cnt.pad, cnt.cull_tests, cnt.warp_iters, cnt.packet_tests, cnt.group_passes, cnt.loop_passes,
                                    cnt.candidates, cnt.shadow_group_passes, cnt.shaded_lights};
#pragma unroll
        for (int k = 0; k < 16; ++k) {
            unsigned long long x = v[k];
            for (int off = 16; off > 0; off >>= 1) x += __shfl_down_sync(kFull, x, off);
            if (lane == 0 && x) atomicAdd(reinterpret_cast<unsigned long long*>(P.counters) + k, x);
        }
    }
}
