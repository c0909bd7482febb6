// rays_roles.cuh — the ray-coherent network kernel with DECOUPLED ROLES (round 2).
//
// Replaces the reference's call sequence main.cu:704-746 (launchSampler -> network->forward ->
// convertHalfToFloat -> launch_volrender_cuda) in render mode, and main.cu:715-728 (network->forward
// on the sampler's rows) in staged mode.
//
// Round 1's kernels (view_fused.cu, network_rays.cu v2) ran every warp through the whole chain
//   encode 16 levels -> barrier -> MMA L0 -> wait -> ReLU -> barrier -> MMA L1 -> wait -> ReLU -> barrier -> MMA L2
// so the ~2000-instruction encoding and the three tensor-core round trips of a tile were serial
// on the same four warps (ncu: issue slots 65 % busy, tensor pipe 6 %).  Here a 512-thread CTA is
//   warps 0-3   the MLP group: thread 0 issues the tcgen05.mma's, the four warps run the TMEM
//               epilogues (ReLU -> fp16 -> A operand of the next layer, written back to TMEM with
//               tcgen05.st: activations never touch shared memory) and, in render mode, warp 0
//               composites the rays front to back (vol_render.cu:55-67);
//   warps 4-15  three encoder groups of four warps.  A tile is 4 sample slots x 32 adjacent rays
//               (lane = ray, warp = slot: one gather instruction touches neighbouring rays at the
//               same depth).  Tiles of the CTA's ray groups are dealt round-robin to the encoder
//               groups; a group writes the 32 features of its 128 rows straight into TMEM as the
//               layer-0 A operand (16 columns, two buffers per group) and never waits for the MLP.
// The MLP group consumes tiles in order.  Producer -> consumer hand-offs are hardware named barriers (encoder
// group -> MLP group per operand buffer, MLP group -> compositor per exchange buffer: rr_named_*), the "buffer is
// free again" signals are mbarriers (enc_free: the MLP warps once layer 0 has read the operand, out_free).
// TMEM per CTA (256 columns, two CTAs per SM): D 64 | A_hidden 32 | 6 x A_encoded 16.
// Shared memory is only the weights (14 KB) and the 4 KB radiance exchange, so ~190 KB of the SM
// stays L1 for the hash-table gathers.
#pragma once
#include "common.cuh"
#include "hashgrid.cuh"
#include "mlp_tile.cuh"

#define RR_MAX_ENC_GROUPS 5                                  // 96 + 2 * 5 * 16 = 256 TMEM columns
#define RR_THREADS(EG) (((EG) + 1) * NET_TILE)
#define RR_CTA_THREADS(EG, RENDER) (RR_THREADS(EG) + ((RENDER) ? 32 : 0))   // render mode: + the compositor warp
#define RR_BUFS(EG) (2 * (EG))
#define RR_COL_D 0
#define RR_COL_AH 64
#define RR_COL_AE 96
#define RR_TMEM_COLS 256
#ifndef RR_GROUPS_PER_CTA
#define RR_GROUPS_PER_CTA 4      // consecutive 32-ray groups one CTA walks (hardware scheduler balances CTAs)
#endif

struct RrSmem {
    static constexpr int W0 = 0;
    static constexpr int W1 = W0 + NET_W * NET_IN * 2;
    static constexpr int W2 = W1 + NET_W * NET_W * 2;
    static constexpr int EXCH = W2 + NET_OUT * NET_W * 2;            // [2][4][32] float4
    static constexpr int BAR = EXCH + 2 * 4 * 32 * 16;     // enc_free[] (+ unused words), mma, out_free[2], tmem slot
    static constexpr int BYTES = BAR + 8 * (2 * RR_BUFS(RR_MAX_ENC_GROUPS) + 5) + 16;
};

struct RrArgs {
    // render mode: sample positions from the CSR entry / exit points, pixels out
    const float* start_points;
    const float* end_points;
    float* pixels;
    // staged mode: sample rows in, radiance / out16 / encoded features out
    const float* samples;
    int stride;
    uint16_t* out16;
    float* radiance4;
    uint16_t* encoded_out;
    // both
    const int* num_hits;
    const int* indices;
    int n_rays, S, out_activation;
    int groups_per_cta;   // consecutive 32-ray groups one CTA walks
    int wait_ticks;   // > 0: mbarrier waits carry this suspend-time hint (development knob)
    int mlp_last;     // 1: the MLP group is the LAST four warps of the CTA (development knob, see rr_launch)
    const float4* seg8;   // optional: packed (sx,sy,sz,ex | ey,ez,0,0) per hit instead of the two arrays / the samples
};

__device__ __forceinline__ void rr_wait(uint64_t* bar, uint32_t parity, int ticks)
{
    if (ticks > 0) tc::mbar_wait_sleep(bar, parity, (uint32_t)ticks);
    else tc::mbar_wait(bar, parity);
}
__device__ __forceinline__ void rr_mlp_barrier() { asm volatile("bar.sync 1, %0;" ::"n"(NET_TILE) : "memory"); }
// Producer -> consumer hand-offs on HARDWARE named barriers: the producers bar.arrive (never block), the
// consumers bar.sync and are descheduled until the count is reached.  The MLP group idles ~85 % of a tile period
// waiting for an encoder group; on an mbarrier all of its 128 threads (and the compositor warp) polled try_wait
// in a loop, ~11 % of the instructions this issue-bound kernel executed (ncu source page, profiles/r02_y).
// ids: 0 __syncthreads, 1 the MLP group, 2 .. 2 + RR_BUFS - 1 operand buffer b (128 encoder + 128 MLP threads),
// then two for the exchange buffers (128 MLP threads + the compositor warp); 14 of the 16 ids at EG = 5.
#define RR_BAR_ENC 2
#define RR_BAR_OUT(EG) (RR_BAR_ENC + RR_BUFS(EG))
__device__ __forceinline__ void rr_named_sync(int id, int count)
{
    asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(count) : "memory");
}
__device__ __forceinline__ void rr_named_arrive(int id, int count)
{
    asm volatile("bar.arrive %0, %1;" ::"r"(id), "r"(count) : "memory");
}

// segment-local sample parameter after i steps of 1/S (sampler.cu:54-66: t += 1.0f/S)
__device__ __forceinline__ float rr_steps(float inv_s, int i, bool pow2)
{
    if (pow2) return rn_mul((float)i, inv_s);   // exact: no rounding ever occurs
    float t = 0.0f;
    for (int k = 0; k < i; k++) t = rn_add(t, inv_s);
    return t;
}

// generic path (configurations the lean paths do not cover): one level through whichever path its mode allows
__device__ __forceinline__ uint32_t rr_encode_generic(const Levels& lv, const uint32_t* __restrict__ tb, int l, float x01,
                                                      float y01, float z01)
{
    const __half2 h = encode_level_any(lv, reinterpret_cast<const char*>(tb), l, x01, y01, z01);
    return *reinterpret_cast<const uint32_t*>(&h);
}

// hidden-layer epilogue of one row: 64 fp32 accumulator columns -> ReLU -> fp16 -> 32 A columns
__device__ __forceinline__ void rr_hidden_epilogue(uint32_t t_d, uint32_t t_a)
{
#pragma unroll
    for (int h = 0; h < NET_W / 32; h++) {
        uint32_t r0[16], r1[16];
        tc::tmem_ld_x16(t_d + h * 32, r0);
        tc::tmem_ld_x16(t_d + h * 32 + 16, r1);
        tc::tmem_ld_wait();
        uint32_t pk[16];
#pragma unroll
        for (int j = 0; j < 8; j++) {
            pk[j] = relu_pack_f16x2(__uint_as_float(r0[2 * j]), __uint_as_float(r0[2 * j + 1]));
            pk[8 + j] = relu_pack_f16x2(__uint_as_float(r1[2 * j]), __uint_as_float(r1[2 * j + 1]));
        }
        tc::tmem_st_x16(t_a + h * 16, pk);
    }
    tc::tmem_st_wait();
}

// one layer with the A operand in TMEM: D[128 x N] = A[128 x K] * W[N x K]^T
template <int N, int K>
__device__ __forceinline__ void rr_issue_layer(uint32_t d_tmem, uint32_t a_tmem, uint32_t w_saddr, uint64_t* bar)
{
    constexpr uint32_t idesc = tc::idesc_f16_f32(NET_TILE, N);
    constexpr uint32_t sbo = (K / 8) * 128;
#pragma unroll
    for (int ks = 0; ks < K / 16; ks++)
        tc::mma_f16_ts(d_tmem, a_tmem + ks * 8, tc::smem_desc(w_saddr + ks * 256, 128, sbo), idesc, ks > 0 ? 1u : 0u);
    tc::mma_commit(bar);
}

// LPI = levels per loop iteration of the encoder (2: two levels' gathers in flight per thread, 1: half the
// loop body, the instruction cache's share of the stalls goes down); LEAN = the lean encoding paths
// (dense levels first, hashed power-of-two levels after), else the generic per-level dispatch.
template <bool RENDER, int EG, int LPI, bool LEAN>
__global__ void __launch_bounds__(RR_CTA_THREADS(EG, RENDER), 2)
rays_roles_kernel(LeanLevels ll, Levels lv, const uint32_t* __restrict__ table32, const uint16_t* __restrict__ weights,
                  RrArgs a)
{
    extern __shared__ __align__(128) uint8_t smem[];
    uint64_t* enc_free = reinterpret_cast<uint64_t*>(smem + RrSmem::BAR);
    uint64_t* mma_bar = enc_free + 2 * RR_BUFS(EG);
    uint64_t* out_free = mma_bar + 3;      // compositor -> MLP group: exch[buf] has been read
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(out_free + 2);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, wk = warp & 3;
    // role index: 0-3 the MLP group, 4 .. 4 * EG + 3 the encoder groups, 4 * EG + 4 the compositor (render mode)
    const int rw = warp >= 4 * (EG + 1) ? warp : (a.mlp_last ? (warp >= 4 * EG ? warp - 4 * EG : warp + 4) : warp);

    load_weight_tile(smem + RrSmem::W0, weights, NET_W, NET_IN, tid, RR_CTA_THREADS(EG, RENDER));
    load_weight_tile(smem + RrSmem::W1, weights + NET_W * NET_IN, NET_W, NET_W, tid, RR_CTA_THREADS(EG, RENDER));
    load_weight_tile(smem + RrSmem::W2, weights + NET_W * NET_IN + NET_W * NET_W, NET_OUT, NET_W, tid, RR_CTA_THREADS(EG, RENDER));
    if (tid == 0) {
        for (int i = 0; i < RR_BUFS(EG); i++) {
            tc::mbar_init(enc_free + i, NET_TILE / 32);   // one arrival per MLP warp (its rows of the operand are read)
        }
        tc::mbar_init(mma_bar, 1);
        for (int i = 0; i < 2; i++) tc::mbar_init(out_free + i, 1);
        tc::fence_mbar_init();
    }
    if (warp == 0) tc::tmem_alloc<RR_TMEM_COLS>(tmem_slot);
    tc::fence_proxy_async_smem();
    tc::fence_before_thread_sync();
    __syncthreads();
    tc::fence_after_thread_sync();
    const uint32_t tmem_base = *tmem_slot;
    const uint32_t lane_sel = (uint32_t)(wk * 32) << 16;     // this warp's TMEM lane quarter
    const uint32_t* tb = table32;
    const int S = a.S;
    const float inv_s = 1.0f / (float)S;
    const bool pow2 = (S & (S - 1)) == 0;
    const int log2s = 31 - __clz(S);
    const int n_groups = (a.n_rays + 31) >> 5;
    const int g_begin = blockIdx.x * a.groups_per_cta;
    const int g_end = min(g_begin + a.groups_per_cta, n_groups);

    if (rw >= 4 && rw < 4 * (EG + 1)) {
        // ------------------------------------------------------------------ encoder group e
        const int e = (rw >> 2) - 1;
        int t_mod = 0;        // (tiles of the CTA before this ray group) mod EG
        int m = 0;            // tiles this group has produced
        for (int g = g_begin; g < g_end; g++) {
            const int ray = g * 32 + lane;
            int ns = 0;
            long long first = 0;
            if (ray < a.n_rays) {
                ns = a.num_hits[ray] * S;
                first = (long long)a.indices[ray];
            }
            int maxn = ns;
#pragma unroll
            for (int o = 16; o; o >>= 1) maxn = max(maxn, __shfl_xor_sync(0xffffffffu, maxn, o));
            const int nt = (maxn + 3) >> 2;
            int t0 = e - t_mod;
            if (t0 < 0) t0 += EG;
            for (int t = t0; t < nt; t += EG, m++) {
                const int b = 2 * e + (m & 1);
                const int q = 4 * t + wk;
                const bool valid = q < ns;
                float x01 = 0.f, y01 = 0.f, z01 = 0.f;
                long long s = 0;
                if (valid) {
                    if (RENDER || !a.samples) {     // (staged mode without a sample buffer: the training forward)
                        // sample position, sampler.cu:45-59
                        const int j = pow2 ? (q >> log2s) : (q / S);
                        const int i = q - j * S;
                        float sx, sy, sz, ex, ey, ez;
                        if (a.seg8) {
                            // one 32-byte sector per lane in ONE request: the lanes' hit lists are ~200 bytes apart, so
                            // each of the six scalar loads below costs the L1 a lookup of 32 different sectors
                            uint32_t rec[8];
                            ldg_v8(a.seg8 + 2 * (first + j), rec);
                            sx = __uint_as_float(rec[0]); sy = __uint_as_float(rec[1]); sz = __uint_as_float(rec[2]);
                            ex = __uint_as_float(rec[3]); ey = __uint_as_float(rec[4]); ez = __uint_as_float(rec[5]);
                        } else {
                            const float* sp = a.start_points + 3 * (first + j);
                            const float* ep = a.end_points + 3 * (first + j);
                            sx = __ldg(sp); sy = __ldg(sp + 1); sz = __ldg(sp + 2);
                            ex = __ldg(ep); ey = __ldg(ep + 1); ez = __ldg(ep + 2);
                        }
                        const float tt = rr_steps(inv_s, i, pow2);
                        x01 = unit_coord(rn_fma(tt, rn_sub(ex, sx), sx));
                        y01 = unit_coord(rn_fma(tt, rn_sub(ey, sy), sy));
                        z01 = unit_coord(rn_fma(tt, rn_sub(ez, sz), sz));
                    } else {
                        s = first * S + q;
                        const float* p = a.samples + s * a.stride;
                        x01 = unit_coord(ld_stream_f1(p));
                        y01 = unit_coord(ld_stream_f1(p + 1));
                        z01 = unit_coord(ld_stream_f1(p + 2));
                    }
                }
                if (m >= 2) {      // the operand buffer is free once layer 0 of its previous tile has completed
                    rr_wait(enc_free + b, (uint32_t)(((m >> 1) - 1) & 1), a.wait_ticks);
                    tc::fence_after_thread_sync();
                }
                const uint32_t t_e = tmem_base + lane_sel + (uint32_t)(RR_COL_AE + 16 * b);
                // feature pair(s) of level(s) l.. -> columns l.. of this row of the operand (and the training copy)
                // (the training copy of the features, encoded_out, is written by the MLP group from this operand:
                // eight 8-byte stores per sample at a 64-byte lane stride were 256 LSU wavefronts per 32 samples on
                // the warps that already keep the LSU pipe 61 % busy with the gathers)
                auto emit1 = [&](int l, uint32_t f) { tc::tmem_st_x1(t_e + l, f); };
                auto emit2 = [&](int l, uint32_t f0, uint32_t f1) { tc::tmem_st_x2(t_e + l, f0, f1); };
                int l = 0;
                if (LEAN) {
                    const int nd = ll.n_dense, nl = ll.n_levels;
                    if (LPI == 2) {
#pragma unroll 1
                        for (; l + 1 < nd; l += 2) {
                            uint32_t f0 = 0u, f1 = 0u;
                            if (valid) {
                                f0 = encode_dense_lean(ll.k[l], tb, x01, y01, z01);
                                f1 = encode_dense_lean(ll.k[l + 1], tb, x01, y01, z01);
                            }
                            emit2(l, f0, f1);
                        }
                        if (l < nd) {     // odd number of dense levels: the boundary pair
                            uint32_t f0 = 0u, f1 = 0u;
                            if (valid) {
                                f0 = encode_dense_lean(ll.k[l], tb, x01, y01, z01);
                                if (l + 1 < nl) f1 = encode_hashed_lean(ll.k[l + 1], tb, x01, y01, z01);
                            }
                            emit2(l, f0, f1);
                            l += 2;
                        }
#pragma unroll 1
                        for (; l + 1 < nl; l += 2) {
                            uint32_t f0 = 0u, f1 = 0u;
                            if (valid) {
                                f0 = encode_hashed_lean(ll.k[l], tb, x01, y01, z01);
                                f1 = encode_hashed_lean(ll.k[l + 1], tb, x01, y01, z01);
                            }
                            emit2(l, f0, f1);
                        }
                        if (l < nl) {
                            uint32_t f0 = 0u;
                            if (valid) f0 = encode_hashed_lean(ll.k[l], tb, x01, y01, z01);
                            emit2(l, f0, 0u);
                            l += 2;
                        }
                    } else {
#pragma unroll 1
                        for (; l < nd; l++) {
                            uint32_t f = 0u;
                            if (valid) f = encode_dense_lean(ll.k[l], tb, x01, y01, z01);
                            emit1(l, f);
                        }
#pragma unroll 1
                        for (; l < nl; l++) {
                            uint32_t f = 0u;
                            if (valid) f = encode_hashed_lean(ll.k[l], tb, x01, y01, z01);
                            emit1(l, f);
                        }
                    }
                } else {
#pragma unroll 1
                    for (; l < lv.n_levels; l++) {
                        uint32_t f = 0u;
                        if (valid) f = rr_encode_generic(lv, tb, l, x01, y01, z01);
                        emit1(l, f);
                    }
                }
#pragma unroll 1
                for (; l < NET_IN / 2; l++) emit1(l, 0u);     // fewer than 16 levels: the rest of the row is zero
                tc::tmem_st_wait();
                tc::fence_before_thread_sync();
                rr_named_arrive(RR_BAR_ENC + b, 2 * NET_TILE);
            }
            t_mod = (t_mod + nt) % EG;
        }
    } else if (rw >= 4 * (EG + 1)) {
        // ------------------------------------------------------------------ compositor warp (render mode)
        // lane = ray.  Composites every tile front to back in the reference's own sequential order
        // (vol_render.cu:55-67: delta carried across segments, tau includes the current sample), off the MLP
        // group's chain: on that chain it was ~220 dependent instructions per tile on one warp while the other
        // three waited at a barrier.
        float4* exch = reinterpret_cast<float4*>(smem + RrSmem::EXCH);
        int T = 0;
        for (int g = g_begin; g < g_end; g++) {
            const int ray = g * 32 + lane;
            int ns = 0;
            if (ray < a.n_rays) ns = a.num_hits[ray] * S;
            int maxn = ns;
#pragma unroll
            for (int o = 16; o; o >>= 1) maxn = max(maxn, __shfl_xor_sync(0xffffffffu, maxn, o));
            const int nt = (maxn + 3) >> 2;
            float tau = 0.0f, t_prev = 0.0f, ax = 0.0f, ay = 0.0f, az = 0.0f;      // vol_render.cu:32-33
            for (int t = 0; t < nt; t++, T++) {
                const int buf = T & 1;
                rr_named_sync(RR_BAR_OUT(EG) + buf, NET_TILE + 32);
#pragma unroll
                for (int k = 0; k < 4; k++) {
                    const int q = 4 * t + k;
                    if (q < ns) {
                        const float4 c = exch[(buf * 4 + k) * 32 + lane];
                        const int i = pow2 ? (q & (S - 1)) : (q % S);
                        const float tv = rn_add(rr_steps(inv_s, i, pow2), inv_s);   // sampler.cu:65-66 t_vals
                        const float delta = fabsf(tv - t_prev);                       // vol_render.cu:56
                        t_prev = tv;
                        const float ds = delta * c.w;
                        tau += ds;                                                    // :60
                        const float w = expf(-tau) * (1.0f - expf(-ds));              // :61-63
                        ax = fmaf(c.x, w, ax);                                        // :65-67
                        ay = fmaf(c.y, w, ay);
                        az = fmaf(c.z, w, az);
                    }
                }
                __syncwarp();
                if (lane == 0) tc::mbar_arrive(out_free + buf);
            }
            if (ray < a.n_rays) {
                float* p = a.pixels + 3 * (long long)ray;     // vol_render.cu:70-72
                p[0] = ax;
                p[1] = ay;
                p[2] = az;
            }
        }
    } else {
        // ------------------------------------------------------------------ MLP group
        const uint32_t t_d = tmem_base + lane_sel + RR_COL_D;
        const uint32_t t_ah = tmem_base + lane_sel + RR_COL_AH;
        const uint32_t s_w0 = tc::smem_u32(smem + RrSmem::W0), s_w1 = tc::smem_u32(smem + RrSmem::W1),
                       s_w2 = tc::smem_u32(smem + RrSmem::W2);
        float4* exch = reinterpret_cast<float4*>(smem + RrSmem::EXCH);
        uint32_t phase = 0;
        int e = 0, m = 0;     // encoder group and per-group tile count of the next tile
        int T = 0;
        for (int g = g_begin; g < g_end; g++) {
            const int ray = g * 32 + lane;
            int ns = 0;
            long long first = 0;
            if (ray < a.n_rays) {
                ns = a.num_hits[ray] * S;
                first = (long long)a.indices[ray];
            }
            int maxn = ns;
#pragma unroll
            for (int o = 16; o; o >>= 1) maxn = max(maxn, __shfl_xor_sync(0xffffffffu, maxn, o));
            const int nt = (maxn + 3) >> 2;
            for (int t = 0; t < nt; t++, T++) {
                const int b = 2 * e + (m & 1);
                // every row of the previous tile's output has been read: D is free
                tc::fence_before_thread_sync();
                rr_named_sync(RR_BAR_ENC + b, 2 * NET_TILE);   // D free AND the operand is there
                if (rw == 0 && tc::elect_one()) {     // the converged issuing warp (tcgen05.cuh: elect_one)
                    tc::fence_after_thread_sync();
                    rr_issue_layer<NET_W, NET_IN>(tmem_base + RR_COL_D, tmem_base + RR_COL_AE + 16 * b, s_w0, mma_bar);
                }
                if (!RENDER && a.encoded_out) {
                    // training copy of this row's 32 features, straight from the layer-0 operand in TMEM (read-only
                    // beside the running MMA): 64 contiguous bytes per sample, four 16-byte stores
                    tc::fence_after_thread_sync();
                    uint32_t f[16];
                    tc::tmem_ld_x16(tmem_base + lane_sel + (uint32_t)(RR_COL_AE + 16 * b), f);
                    tc::tmem_ld_wait();
                    const int q = 4 * t + wk;
                    if (q < ns) {
                        uint16_t* o = a.encoded_out + (first * S + q) * NET_IN;        // 64-byte rows
                        const uint32_t lo[8] = {f[0], f[1], f[2], f[3], f[4], f[5], f[6], f[7]};
                        const uint32_t hi[8] = {f[8], f[9], f[10], f[11], f[12], f[13], f[14], f[15]};
                        stg_v8(o, lo);
                        stg_v8(o + 16, hi);
                    }
                }
                rr_wait(mma_bar, phase, a.wait_ticks);
                phase ^= 1;
                tc::fence_after_thread_sync();
                __syncwarp();
                if (lane == 0) tc::mbar_arrive(enc_free + b);
                rr_hidden_epilogue(t_d, t_ah);
                tc::fence_before_thread_sync();
                rr_mlp_barrier();
                if (rw == 0 && tc::elect_one()) {     // the converged issuing warp (tcgen05.cuh: elect_one)
                    tc::fence_after_thread_sync();
                    rr_issue_layer<NET_W, NET_W>(tmem_base + RR_COL_D, tmem_base + RR_COL_AH, s_w1, mma_bar);
                }
                rr_wait(mma_bar, phase, a.wait_ticks);
                phase ^= 1;
                tc::fence_after_thread_sync();
                rr_hidden_epilogue(t_d, t_ah);
                tc::fence_before_thread_sync();
                rr_mlp_barrier();
                if (rw == 0 && tc::elect_one()) {     // the converged issuing warp (tcgen05.cuh: elect_one)
                    tc::fence_after_thread_sync();
                    rr_issue_layer<NET_OUT, NET_W>(tmem_base + RR_COL_D, tmem_base + RR_COL_AH, s_w2, mma_bar);
                }
                rr_wait(mma_bar, phase, a.wait_ticks);
                phase ^= 1;
                tc::fence_after_thread_sync();
                uint32_t r[16];
                output_row_load(t_d, r);
                if (RENDER) {
                    const int buf = T & 1;
                    if (T >= 2) rr_wait(out_free + buf, (uint32_t)(((T >> 1) - 1) & 1), a.wait_ticks);
                    exch[(buf * 4 + wk) * 32 + lane] = output_row_rgba(r, a.out_activation);
                    rr_named_arrive(RR_BAR_OUT(EG) + buf, NET_TILE + 32);
                } else {
                    const int q = 4 * t + wk;
                    if (q < ns) output_row_store(r, a.out_activation, first * S + q, a.out16, a.radiance4);
                }
                if (++e == EG) {
                    e = 0;
                    m++;
                }
            }
        }
    }
    tc::fence_before_thread_sync();
    __syncthreads();
    tc::fence_after_thread_sync();
    if (warp == 0) tc::tmem_dealloc<RR_TMEM_COLS>(tmem_base);
}

// launch helper shared by the two entry points (render_fused.cu, network_rays.cu)
int rr_launch(bool render, const rtxnerf_hashgrid* enc, const rtxnerf_mlp* mlp, const Levels& lv, const RrArgs& a,
              cudaStream_t stream);
