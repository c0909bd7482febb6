// view_fused.cu — sampler + hash encoding + fused MLP + alpha compositing in ONE kernel.
//
// The render path of the reference's per-ray pipeline after the march:
//   launchSampler        sampler/sampler.cu:14-131      (REGULAR mode)
//   network->forward     main.cu:715-728                (hash-grid encoding + 2x64 MLP, see network_rays.cu)
//   launch_volrender     vol_render/vol_render.cu:19-73,144-164
// The staged entry points (rtxnerf_launch_sampler, rtxnerf_network_forward_rays,
// rtxnerf_launch_volrender) keep the reference's buffers: 20 B of sample + 4 B of t written and
// read back, 16 B of radiance written and read back: 64 B of HBM traffic per sample that carries
// no information the march output does not already hold.  Here a sample exists only in
// registers: position from the CSR entry/exit points (sampler.cu:45-59), features and layers as
// in network_forward_rays2_kernel, and the radiance goes through 4 KB of shared memory to the
// warp that composites the ray front to back in the reference's own sequential order
// (vol_render.cu:55-67: delta carried across segments, tau includes the current sample).
// Output: pixels[n_rays][3] only.  Training keeps the staged path (its backward needs the
// buffers).
#include <stdlib.h>

#include "common.cuh"
#include "hashgrid.cuh"
#include "mlp_tile.cuh"

#define VF_WGS 4
#define VF_THREADS (VF_WGS * NET_TILE)

struct VfSmem {
    static constexpr int W0 = 0;
    static constexpr int W1 = W0 + NET_W * NET_IN * 2;
    static constexpr int W2 = W1 + NET_W * NET_W * 2;
    static constexpr int ACT = W2 + NET_OUT * NET_W * 2;                  // VF_WGS tiles of 128 x 64 fp16
    static constexpr int EXCH = ACT + VF_WGS * NET_TILE * NET_W * 2;      // [VF_WGS][2][4][32] float4
    static constexpr int BAR = EXCH + VF_WGS * 2 * 4 * 32 * 16;           // VF_WGS mbarriers, tmem slot
    static constexpr int BYTES = BAR + 8 * VF_WGS + 16;
};

__device__ __forceinline__ void vf_wg_barrier(int wg)
{
    asm volatile("bar.sync %0, %1;" ::"r"(wg + 1), "n"(NET_TILE) : "memory");
}

// Two levels at a time through the straight-line fast paths (same dispatch as network_rays.cu).
__device__ __forceinline__ void vf_encode_pair(const Levels& lv, const char* __restrict__ tb, int l, float x01,
                                               float y01, float z01, uint32_t& f0, uint32_t& f1)
{
    __half2 h0 = __floats2half2_rn(0.f, 0.f), h1 = h0;
    const bool two = l + 1 < lv.n_levels;
    const uint32_t m0 = lv.mode[l], m1 = two ? lv.mode[l + 1] : 2u;
    if (m0 == 1u && m1 == 1u) {
        h0 = encode_level_fast<1>(lv, l, tb, x01, y01, z01);
        h1 = encode_level_fast<1>(lv, l + 1, tb, x01, y01, z01);
    } else if (m0 == 0u && m1 == 0u) {
        h0 = encode_level_fast<0>(lv, l, tb, x01, y01, z01);
        h1 = encode_level_fast<0>(lv, l + 1, tb, x01, y01, z01);
    } else if (m0 == 0u && m1 == 1u) {
        h0 = encode_level_fast<0>(lv, l, tb, x01, y01, z01);
        h1 = encode_level_fast<1>(lv, l + 1, tb, x01, y01, z01);
    } else {
        if (l < lv.n_levels) h0 = encode_level_any(lv, tb, l, x01, y01, z01);
        if (two) h1 = encode_level_any(lv, tb, l + 1, x01, y01, z01);
    }
    f0 = *reinterpret_cast<const uint32_t*>(&h0);
    f1 = *reinterpret_cast<const uint32_t*>(&h1);
}

// segment-local sample parameter after i steps of 1/S (sampler.cu:54-66: t += 1.0f/S)
__device__ __forceinline__ float vf_steps(float inv_s, int i, bool pow2)
{
    if (pow2) return rn_mul((float)i, inv_s);   // exact: no rounding ever occurs
    float t = 0.0f;
    for (int k = 0; k < i; k++) t = rn_add(t, inv_s);
    return t;
}

__global__ void __launch_bounds__(VF_THREADS, 2)
view_fused_kernel(Levels lv, const uint32_t* __restrict__ table32, const uint16_t* __restrict__ weights,
                  int out_activation, const float* __restrict__ start_points, const float* __restrict__ end_points,
                  const int* __restrict__ num_hits, const int* __restrict__ indices, int n_rays, int S,
                  float* __restrict__ pixels)
{
    extern __shared__ __align__(128) uint8_t smem[];
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + VfSmem::BAR);
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + VfSmem::BAR + 8 * VF_WGS);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int wg = tid >> 7, wt = tid & (NET_TILE - 1), wk = warp & 3;

    load_weight_tile(smem + VfSmem::W0, weights, NET_W, NET_IN, tid, VF_THREADS);
    load_weight_tile(smem + VfSmem::W1, weights + NET_W * NET_IN, NET_W, NET_W, tid, VF_THREADS);
    load_weight_tile(smem + VfSmem::W2, weights + NET_W * NET_IN + NET_W * NET_W, NET_OUT, NET_W, tid, VF_THREADS);
    if (tid == 0) {
        for (int i = 0; i < VF_WGS; i++) tc::mbar_init(bars + i, 1);
        tc::fence_mbar_init();
    }
    if (warp == 0) tc::tmem_alloc<VF_WGS * NET_TMEM_COLS>(tmem_slot);
    tc::fence_proxy_async_smem();
    tc::fence_before_thread_sync();
    __syncthreads();
    tc::fence_after_thread_sync();
    const uint32_t tmem_wg = *tmem_slot + (uint32_t)(wg * NET_TMEM_COLS);
    const uint32_t t_row = tmem_wg + ((uint32_t)(wk * 32) << 16);
    uint8_t* act = smem + VfSmem::ACT + wg * (NET_TILE * NET_W * 2);
    float4* exch = reinterpret_cast<float4*>(smem + VfSmem::EXCH) + wg * (2 * 4 * 32);
    const uint32_t s_w0 = tc::smem_u32(smem + VfSmem::W0), s_w1 = tc::smem_u32(smem + VfSmem::W1),
                   s_w2 = tc::smem_u32(smem + VfSmem::W2), s_act = tc::smem_u32(act);
    const char* tb = reinterpret_cast<const char*>(table32);
    uint64_t* bar = bars + wg;
    uint32_t phase = 0;
    const float inv_s = 1.0f / (float)S;
    const bool pow2 = (S & (S - 1)) == 0;
    const int log2s = 31 - __clz(S);

    const int n_groups = (n_rays + 31) >> 5;
    for (int g = blockIdx.x * VF_WGS + wg; g < n_groups; g += gridDim.x * VF_WGS) {
        const int ray = g * 32 + lane;
        int ns = 0;
        long long first = 0;
        if (ray < n_rays) {
            ns = num_hits[ray] * S;
            first = (long long)indices[ray];
        }
        int maxn = ns;
#pragma unroll
        for (int o = 16; o; o >>= 1) maxn = max(maxn, __shfl_xor_sync(0xffffffffu, maxn, o));

        // compositing state of this lane's ray (used by warp 0 of the group), vol_render.cu:32-33
        float tau = 0.0f, t_prev = 0.0f, ax = 0.0f, ay = 0.0f, az = 0.0f;
        int pend_q0 = -1, it = 0;
        // composites the four samples q0..q0+3 parked in exch[buf] (front to back)
        auto composite4 = [&](int buf, int q0) {
#pragma unroll
            for (int k = 0; k < 4; k++) {
                const int q = q0 + k;
                if (q < ns) {
                    const float4 c = exch[(buf * 4 + k) * 32 + lane];
                    const int i = pow2 ? (q & (S - 1)) : (q % S);
                    const float tv = rn_add(vf_steps(inv_s, i, pow2), inv_s);   // sampler.cu:65-66 t_vals
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
        };

        for (int q0 = 0; q0 < maxn; q0 += 4, it++) {
            const int q = q0 + wk;
            const bool valid = q < ns;
            // ---- sample position (sampler.cu:45-59) and its 16 feature pairs -> this row of the A tile
            if (valid) {
                const int j = pow2 ? (q >> log2s) : (q / S);
                const int i = q - j * S;
                const float* a = start_points + 3 * (first + j);
                const float* b = end_points + 3 * (first + j);
                const float sx = __ldg(a), sy = __ldg(a + 1), sz = __ldg(a + 2);
                const float t = vf_steps(inv_s, i, pow2);
                const float px = rn_fma(t, rn_sub(__ldg(b), sx), sx);
                const float py = rn_fma(t, rn_sub(__ldg(b + 1), sy), sy);
                const float pz = rn_fma(t, rn_sub(__ldg(b + 2), sz), sz);
                const float x01 = unit_coord(px), y01 = unit_coord(py), z01 = unit_coord(pz);
#pragma unroll 1
                for (int c = 0; c < NET_IN / 8; c++) {
                    uint4 f;
                    vf_encode_pair(lv, tb, 4 * c, x01, y01, z01, f.x, f.y);
                    vf_encode_pair(lv, tb, 4 * c + 2, x01, y01, z01, f.z, f.w);
                    *reinterpret_cast<uint4*>(act + tile_chunk_off(wt, c, NET_IN)) = f;
                }
            } else {
#pragma unroll
                for (int c = 0; c < NET_IN / 8; c++)
                    *reinterpret_cast<uint4*>(act + tile_chunk_off(wt, c, NET_IN)) = make_uint4(0u, 0u, 0u, 0u);
            }
            tc::fence_proxy_async_smem();
            tc::fence_before_thread_sync();
            vf_wg_barrier(wg);
            if (wt == 0) {
                tc::fence_after_thread_sync();
                issue_layer<NET_W, NET_IN>(tmem_wg, s_act, s_w0, bar);
            }
            // the previous iteration's radiance is complete behind the barrier: composite it
            // while the tensor core runs layer 0
            if (wk == 0 && pend_q0 >= 0) composite4((it - 1) & 1, pend_q0);
            tc::mbar_wait(bar, phase);
            phase ^= 1;
            tc::fence_after_thread_sync();
            relu_row_to_smem(t_row, act, wt);
            tc::fence_proxy_async_smem();
            tc::fence_before_thread_sync();
            vf_wg_barrier(wg);
            if (wt == 0) {
                tc::fence_after_thread_sync();
                issue_layer<NET_W, NET_W>(tmem_wg, s_act, s_w1, bar);
            }
            tc::mbar_wait(bar, phase);
            phase ^= 1;
            tc::fence_after_thread_sync();
            relu_row_to_smem(t_row, act, wt);
            tc::fence_proxy_async_smem();
            tc::fence_before_thread_sync();
            vf_wg_barrier(wg);
            if (wt == 0) {
                tc::fence_after_thread_sync();
                issue_layer<NET_OUT, NET_W>(tmem_wg, s_act, s_w2, bar);
            }
            tc::mbar_wait(bar, phase);
            phase ^= 1;
            tc::fence_after_thread_sync();
            uint32_t r[16];
            output_row_load(t_row, r);
            tc::fence_before_thread_sync();
            exch[((it & 1) * 4 + wk) * 32 + lane] = output_row_rgba(r, out_activation);
            pend_q0 = q0;
        }
        if (maxn > 0) {
            vf_wg_barrier(wg);
            if (wk == 0) composite4((it - 1) & 1, pend_q0);
        }
        if (wk == 0 && ray < n_rays) {
            float* p = pixels + 3 * (long long)ray;     // vol_render.cu:70-72
            p[0] = ax;
            p[1] = ay;
            p[2] = az;
        }
    }
    tc::fence_before_thread_sync();
    __syncthreads();
    tc::fence_after_thread_sync();
    if (warp == 0) tc::tmem_dealloc<VF_WGS * NET_TMEM_COLS>(*tmem_slot);
}

// Round-1 kernel, kept for A/B measurements (RTXNERF_FUSED_V1=1); the product entry point is in
// render_fused.cu.
int vf1_render_rays_fused(const rtxnerf_hashgrid* enc, const rtxnerf_mlp* mlp, const float* d_start_points,
                          const float* d_end_points, const int* d_num_hits, const int* d_indices, int n_rays,
                          int samples_per_segment, int sample_type, float* d_pixels, void* stream)
{
    if (!enc || !mlp || !mlp->weights || !d_start_points || !d_end_points || !d_num_hits || !d_indices || !d_pixels)
        return RTXNERF_EINVAL;
    if (n_rays < 0 || samples_per_segment < 1) return RTXNERF_EINVAL;
    if (sample_type != 0) return RTXNERF_ECONFIG;   // REGULAR only (the reference's parity mode)
    if (mlp->in_dim != NET_IN || mlp->width != NET_W || mlp->n_hidden != 2) return RTXNERF_ECONFIG;
    if (mlp->out_activation != 0 && mlp->out_activation != 1) return RTXNERF_ECONFIG;
    Levels lv;
    int rc = levels_from(enc, &lv);
    if (rc) return rc;
    if (enc->n_levels * 2 > NET_IN) return RTXNERF_ECONFIG;
    if (n_rays == 0) return RTXNERF_OK;
    static RtxPerDeviceOnce attr;
    RTX_CUDA_TRY((cudaError_t)rtx_set_max_smem(attr, view_fused_kernel, VfSmem::BYTES));
    const long long groups = ((long long)n_rays + 31) / 32;
    long long blocks = (groups + VF_WGS - 1) / VF_WGS;
    // Not a resident-only grid: one CTA per VF_WGS groups (up to 64 CTAs per SM, 2 resident) lets the hardware
    // scheduler balance the groups.  The static grid-stride split over 2 CTAs per SM measured 5 % slower on the
    // 800 x 800 view (4.63 -> 4.39 ms); the per-CTA set-up (14 KB of weights, TMEM allocation) is noise.
#ifndef VF_BLOCKS_PER_SM
#define VF_BLOCKS_PER_SM 64
#endif
    if (blocks > (long long)RTX_NUM_SMS * VF_BLOCKS_PER_SM) blocks = (long long)RTX_NUM_SMS * VF_BLOCKS_PER_SM;
    view_fused_kernel<<<(unsigned)blocks, VF_THREADS, VfSmem::BYTES, (cudaStream_t)stream>>>(
        lv, reinterpret_cast<const uint32_t*>(enc->table), mlp->weights, mlp->out_activation, d_start_points,
        d_end_points, d_num_hits, d_indices, n_rays, samples_per_segment, d_pixels);
    return rtx_launch_status();
}
