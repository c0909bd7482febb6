// network_rays.cu — ray-coherent hash encoding + fused MLP forward for sm_100a.
//
// Same function as network.cu's flat kernel (tiny-cuda-nn encoding + FullyFusedMLP forward as
// called at main.cu:715-728) over the sampler's CSR sample buffer, but scheduled by RAY:
//
//   * a warp-group (4 warps) owns 32 consecutive rays (adjacent pixels of an image row);
//     lane = ray, warp k = sample slot q0+k along the ray.  One gather instruction therefore
//     touches the 32 neighbouring rays at the SAME depth, whose corners fall into the same or
//     adjacent cells on almost every level: the number of distinct 128-byte lines per load
//     (the L1TEX wavefronts that bound the flat kernel, profiles/r01_b_*) drops from up to 32
//     to a handful, and the four warps walk consecutive samples so coarse levels hit L1.
//   * the 16 feature pairs of a row stay in registers while they are gathered; the CTA's single
//     16 KB activation tile and 64 TMEM columns are taken only for the ~1.5k cycles of the three
//     tcgen05 layers (a shared-memory mutex between the CTA's 4 warp-groups).  30 KB of shared
//     memory per 512 threads leaves ~160 KB of the SM's L1 to the hash table.
//   * weights stay resident in shared memory (UMMA K-major layout) for the CTA's lifetime.
#include <stdlib.h>

#include "common.cuh"
#include "hashgrid.cuh"
#include "mlp_tile.cuh"
#include "rays_roles.cuh"

#define NETR_WGS 4
#define NETR_THREADS (NETR_WGS * NET_TILE)

struct NetRSmem {
    static constexpr int W0 = 0;
    static constexpr int W1 = W0 + NET_W * NET_IN * 2;
    static constexpr int W2 = W1 + NET_W * NET_W * 2;
    static constexpr int ACT = W2 + NET_OUT * NET_W * 2;
    static constexpr int BAR = ACT + NET_TILE * NET_W * 2;   // NETR_WGS mbarriers, tmem slot, lock
    static constexpr int BYTES = BAR + 8 * NETR_WGS + 16;
};

__device__ __forceinline__ void wg_barrier(int wg)
{
    asm volatile("bar.sync %0, %1;" ::"r"(wg + 1), "n"(NET_TILE) : "memory");
}

// Two levels at a time.  When both take the same straight-line fast path (the common case:
// coarse levels dense, fine levels hashed) the gathers of both are in flight together.
__device__ __forceinline__ void encode_pair(const Levels& lv, const uint32_t* __restrict__ table32, int l,
                                            float x01, float y01, float z01, uint32_t& f0, uint32_t& f1)
{
    const char* tb = reinterpret_cast<const char*>(table32);
    __half2 h0 = __floats2half2_rn(0.f, 0.f), h1 = h0;
    const bool two = l + 1 < lv.n_levels;
    const uint32_t m0 = lv.mode[l], m1 = two ? lv.mode[l + 1] : 2u;
    if (m0 == 1u && m1 == 1u) {
        h0 = encode_level_fast<1>(lv, l, tb, x01, y01, z01);
        h1 = encode_level_fast<1>(lv, l + 1, tb, x01, y01, z01);
    } else if (m0 == 0u && m1 == 0u) {
        h0 = encode_level_fast<0>(lv, l, tb, x01, y01, z01);
        h1 = encode_level_fast<0>(lv, l + 1, tb, x01, y01, z01);
    } else if (m0 == 0u && m1 == 1u) {      // the pair that straddles the dense -> hashed boundary
        h0 = encode_level_fast<0>(lv, l, tb, x01, y01, z01);
        h1 = encode_level_fast<1>(lv, l + 1, tb, x01, y01, z01);
    } else {
        if (l < lv.n_levels) h0 = encode_level_any(lv, tb, l, x01, y01, z01);
        if (two) h1 = encode_level_any(lv, tb, l + 1, x01, y01, z01);
    }
    f0 = *reinterpret_cast<const uint32_t*>(&h0);
    f1 = *reinterpret_cast<const uint32_t*>(&h1);
}

template <int MIN_CTAS>
__global__ void __launch_bounds__(NETR_THREADS, MIN_CTAS)
network_forward_rays_kernel(Levels lv, const uint32_t* __restrict__ table32, const uint16_t* __restrict__ weights,
                            const float* __restrict__ samples, int stride, int out_activation,
                            const int* __restrict__ num_hits, const int* __restrict__ indices, int n_rays, int S,
                            uint16_t* __restrict__ out16, float* __restrict__ radiance4,
                            uint16_t* __restrict__ encoded_out)
{
    extern __shared__ __align__(128) uint8_t smem[];
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + NetRSmem::BAR);
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + NetRSmem::BAR + 8 * NETR_WGS);
    int* lock = reinterpret_cast<int*>(smem + NetRSmem::BAR + 8 * NETR_WGS + 4);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int wg = tid >> 7, wt = tid & (NET_TILE - 1), wk = warp & 3;

    load_weight_tile(smem + NetRSmem::W0, weights, NET_W, NET_IN, tid, NETR_THREADS);
    load_weight_tile(smem + NetRSmem::W1, weights + NET_W * NET_IN, NET_W, NET_W, tid, NETR_THREADS);
    load_weight_tile(smem + NetRSmem::W2, weights + NET_W * NET_IN + NET_W * NET_W, NET_OUT, NET_W, tid, NETR_THREADS);
    if (tid == 0) {
        for (int i = 0; i < NETR_WGS; i++) tc::mbar_init(bars + i, 1);
        *lock = 0;
        tc::fence_mbar_init();
    }
    if (warp == 0) tc::tmem_alloc<NET_TMEM_COLS>(tmem_slot);
    tc::fence_proxy_async_smem();
    tc::fence_before_thread_sync();
    __syncthreads();
    tc::fence_after_thread_sync();
    const uint32_t tmem_base = *tmem_slot;
    const uint32_t t_row = tmem_base + ((uint32_t)(wk * 32) << 16);   // TMEM lanes of this warp's quarter
    const uint32_t s_w0 = tc::smem_u32(smem + NetRSmem::W0), s_w1 = tc::smem_u32(smem + NetRSmem::W1),
                   s_w2 = tc::smem_u32(smem + NetRSmem::W2), s_act = tc::smem_u32(smem + NetRSmem::ACT);
    uint8_t* act = smem + NetRSmem::ACT;
    uint64_t* bar = bars + wg;
    uint32_t phase = 0;

    const int n_groups = (n_rays + 31) >> 5;
    for (int g = blockIdx.x * NETR_WGS + wg; g < n_groups; g += gridDim.x * NETR_WGS) {
        const int ray = g * 32 + lane;
        int ns = 0;
        long long base = 0;
        if (ray < n_rays) {
            ns = num_hits[ray] * S;
            base = (long long)indices[ray] * S;
        }
        int maxn = ns;
#pragma unroll
        for (int o = 16; o; o >>= 1) maxn = max(maxn, __shfl_xor_sync(0xffffffffu, maxn, o));

        for (int q0 = 0; q0 < maxn; q0 += 4) {
            const int q = q0 + wk;
            const bool valid = q < ns;
            const long long s = base + q;
            // ---- gather: this row's 16 feature pairs, in registers
            uint32_t pk[NET_IN / 2];
#pragma unroll
            for (int i = 0; i < NET_IN / 2; i++) pk[i] = 0u;
            if (valid) {
                const float* p = samples + s * stride;
                const float x01 = unit_coord(ld_stream_f1(p));
                const float y01 = unit_coord(ld_stream_f1(p + 1));
                const float z01 = unit_coord(ld_stream_f1(p + 2));
#pragma unroll 1
                for (int c = 0; c < NET_IN / 4; c++) {
                    uint32_t f0, f1;
                    encode_pair(lv, table32, 2 * c, x01, y01, z01, f0, f1);
#pragma unroll
                    for (int i = 0; i < NET_IN / 4; i++)
                        if (i == c) {
                            pk[2 * i] = f0;
                            pk[2 * i + 1] = f1;
                        }
                }
            }
            if (encoded_out && valid) {   // training forward: keep the features for the backward pass
                uint4* eo = reinterpret_cast<uint4*>(encoded_out + s * NET_IN);
#pragma unroll
                for (int c = 0; c < NET_IN / 8; c++)
                    eo[c] = make_uint4(pk[4 * c], pk[4 * c + 1], pk[4 * c + 2], pk[4 * c + 3]);
            }
            // ---- the three layers, holding the CTA's activation tile and TMEM columns
            if (wt == 0) {
                while (atomicCAS(lock, 0, 1) != 0) __nanosleep(64);
                __threadfence_block();
            }
            wg_barrier(wg);
#pragma unroll
            for (int c = 0; c < NET_IN / 8; c++)
                *reinterpret_cast<uint4*>(act + tile_chunk_off(wt, c, NET_IN)) =
                    make_uint4(pk[4 * c], pk[4 * c + 1], pk[4 * c + 2], pk[4 * c + 3]);
            tc::fence_proxy_async_smem();
            tc::fence_before_thread_sync();
            wg_barrier(wg);
            if (wt == 0) {
                tc::fence_after_thread_sync();
                issue_layer<NET_W, NET_IN>(tmem_base, s_act, s_w0, bar);
            }
            tc::mbar_wait(bar, phase);
            phase ^= 1;
            tc::fence_after_thread_sync();
            relu_row_to_smem(t_row, act, wt);
            tc::fence_proxy_async_smem();
            tc::fence_before_thread_sync();
            wg_barrier(wg);
            if (wt == 0) {
                tc::fence_after_thread_sync();
                issue_layer<NET_W, NET_W>(tmem_base, s_act, s_w1, bar);
            }
            tc::mbar_wait(bar, phase);
            phase ^= 1;
            tc::fence_after_thread_sync();
            relu_row_to_smem(t_row, act, wt);
            tc::fence_proxy_async_smem();
            tc::fence_before_thread_sync();
            wg_barrier(wg);
            if (wt == 0) {
                tc::fence_after_thread_sync();
                issue_layer<NET_OUT, NET_W>(tmem_base, s_act, s_w2, bar);
            }
            tc::mbar_wait(bar, phase);
            phase ^= 1;
            tc::fence_after_thread_sync();
            uint32_t r[16];
            output_row_load(t_row, r);
            tc::fence_before_thread_sync();
            wg_barrier(wg);                      // every row of the accumulator has been read
            if (wt == 0) {
                __threadfence_block();
                atomicExch(lock, 0);
            }
            if (valid) output_row_store(r, out_activation, s, out16, radiance4);
        }
    }
    tc::fence_before_thread_sync();
    __syncthreads();
    tc::fence_after_thread_sync();
    if (warp == 0) tc::tmem_dealloc<NET_TMEM_COLS>(tmem_base);
}

// ---------------------------------------------------------------------------------------
// Version 2: no mutex.  Every warp-group owns a 16 KB activation tile and 64 TMEM columns
// (4 x 64 = 256 columns per CTA, two CTAs fill the SM's 512), so the four groups of a CTA
// never wait on each other: the tensor pipe interleaves their layers.  Features go from the
// blend straight into the group's A tile (one 16-byte store per four levels) instead of
// being held in 16 registers until the tile is free; those registers go to gathers in flight.
// ---------------------------------------------------------------------------------------
template <int WGS>
struct NetR2Smem {
    static constexpr int W0 = 0;
    static constexpr int W1 = W0 + NET_W * NET_IN * 2;
    static constexpr int W2 = W1 + NET_W * NET_W * 2;
    static constexpr int ACT = W2 + NET_OUT * NET_W * 2;                 // WGS tiles of 128 x 64 fp16
    static constexpr int BAR = ACT + WGS * NET_TILE * NET_W * 2;    // WGS mbarriers, tmem slot
    static constexpr int BYTES = BAR + 8 * WGS + 16;
};

template <int WGS, int MINB>
__global__ void __launch_bounds__(WGS * NET_TILE, MINB)
network_forward_rays2_kernel(Levels lv, const uint32_t* __restrict__ table32, const uint16_t* __restrict__ weights,
                             const float* __restrict__ samples, int stride, int out_activation,
                             const int* __restrict__ num_hits, const int* __restrict__ indices, int n_rays, int S,
                             uint16_t* __restrict__ out16, float* __restrict__ radiance4,
                             uint16_t* __restrict__ encoded_out)
{
    extern __shared__ __align__(128) uint8_t smem[];
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + NetR2Smem<WGS>::BAR);
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + NetR2Smem<WGS>::BAR + 8 * WGS);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int wg = tid >> 7, wt = tid & (NET_TILE - 1), wk = warp & 3;

    load_weight_tile(smem + NetR2Smem<WGS>::W0, weights, NET_W, NET_IN, tid, (WGS * NET_TILE));
    load_weight_tile(smem + NetR2Smem<WGS>::W1, weights + NET_W * NET_IN, NET_W, NET_W, tid, (WGS * NET_TILE));
    load_weight_tile(smem + NetR2Smem<WGS>::W2, weights + NET_W * NET_IN + NET_W * NET_W, NET_OUT, NET_W, tid, (WGS * NET_TILE));
    if (tid == 0) {
        for (int i = 0; i < WGS; i++) tc::mbar_init(bars + i, 1);
        tc::fence_mbar_init();
    }
    if (warp == 0) tc::tmem_alloc<(WGS == 3 ? 4 : WGS) * NET_TMEM_COLS>(tmem_slot);
    tc::fence_proxy_async_smem();
    tc::fence_before_thread_sync();
    __syncthreads();
    tc::fence_after_thread_sync();
    const uint32_t tmem_wg = *tmem_slot + (uint32_t)(wg * NET_TMEM_COLS);
    const uint32_t t_row = tmem_wg + ((uint32_t)(wk * 32) << 16);   // TMEM lanes of this warp's quarter
    uint8_t* act = smem + NetR2Smem<WGS>::ACT + wg * (NET_TILE * NET_W * 2);
    const uint32_t s_w0 = tc::smem_u32(smem + NetR2Smem<WGS>::W0), s_w1 = tc::smem_u32(smem + NetR2Smem<WGS>::W1),
                   s_w2 = tc::smem_u32(smem + NetR2Smem<WGS>::W2), s_act = tc::smem_u32(act);
    uint64_t* bar = bars + wg;
    uint32_t phase = 0;

    const int n_groups = (n_rays + 31) >> 5;
    for (int g = blockIdx.x * WGS + wg; g < n_groups; g += gridDim.x * WGS) {
        const int ray = g * 32 + lane;
        int ns = 0;
        long long base = 0;
        if (ray < n_rays) {
            ns = num_hits[ray] * S;
            base = (long long)indices[ray] * S;
        }
        int maxn = ns;
#pragma unroll
        for (int o = 16; o; o >>= 1) maxn = max(maxn, __shfl_xor_sync(0xffffffffu, maxn, o));

        for (int q0 = 0; q0 < maxn; q0 += 4) {
            const int q = q0 + wk;
            const bool valid = q < ns;
            const long long s = base + q;
            // ---- gather: four levels (one 16-byte chunk of this row of the A tile) at a time
            if (valid) {
                const float* p = samples + s * stride;
                const float x01 = unit_coord(ld_stream_f1(p));
                const float y01 = unit_coord(ld_stream_f1(p + 1));
                const float z01 = unit_coord(ld_stream_f1(p + 2));
#pragma unroll 1
                for (int c = 0; c < NET_IN / 8; c++) {
                    uint4 f;
                    encode_pair(lv, table32, 4 * c, x01, y01, z01, f.x, f.y);
                    encode_pair(lv, table32, 4 * c + 2, x01, y01, z01, f.z, f.w);
                    *reinterpret_cast<uint4*>(act + tile_chunk_off(wt, c, NET_IN)) = f;
                    if (encoded_out) reinterpret_cast<uint4*>(encoded_out + s * NET_IN)[c] = f;
                }
            } else {
#pragma unroll
                for (int c = 0; c < NET_IN / 8; c++)
                    *reinterpret_cast<uint4*>(act + tile_chunk_off(wt, c, NET_IN)) = make_uint4(0u, 0u, 0u, 0u);
            }
            tc::fence_proxy_async_smem();
            tc::fence_before_thread_sync();
            wg_barrier(wg);
            if (wt == 0) {
                tc::fence_after_thread_sync();
                issue_layer<NET_W, NET_IN>(tmem_wg, s_act, s_w0, bar);
            }
            tc::mbar_wait(bar, phase);
            phase ^= 1;
            tc::fence_after_thread_sync();
            relu_row_to_smem(t_row, act, wt);
            tc::fence_proxy_async_smem();
            tc::fence_before_thread_sync();
            wg_barrier(wg);
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
            wg_barrier(wg);
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
            if (valid) output_row_store(r, out_activation, s, out16, radiance4);
        }
    }
    tc::fence_before_thread_sync();
    __syncthreads();
    tc::fence_after_thread_sync();
    if (warp == 0) tc::tmem_dealloc<(WGS == 3 ? 4 : WGS) * NET_TMEM_COLS>(*tmem_slot);
}

extern "C" int rtxnerf_network_forward_rays(const rtxnerf_hashgrid* enc, const rtxnerf_mlp* mlp,
                                            const float* d_samples, int sample_stride, const int* d_num_hits,
                                            const int* d_indices, int n_rays, int samples_per_segment,
                                            uint16_t* d_out16, float* d_radiance4, uint16_t* d_encoded_out,
                                            void* stream)
{
    if (!enc || !mlp || !mlp->weights || !d_samples || !d_num_hits || !d_indices || (!d_out16 && !d_radiance4))
        return RTXNERF_EINVAL;
    if (n_rays < 0 || samples_per_segment < 1 || sample_stride < 3) return RTXNERF_EINVAL;
    if ((uintptr_t)d_encoded_out & 31u) return RTXNERF_EINVAL;      // 64-byte rows written with 256-bit stores
    if (mlp->in_dim != NET_IN || mlp->width != NET_W || mlp->n_hidden != 2) return RTXNERF_ECONFIG;
    if (mlp->out_activation != 0 && mlp->out_activation != 1) return RTXNERF_ECONFIG;
    Levels lv;
    int rc = levels_from(enc, &lv);
    if (rc) return rc;
    if (enc->n_levels * 2 > NET_IN) return RTXNERF_ECONFIG;
    if (n_rays == 0) return RTXNERF_OK;
    // development knobs (read once per process): RTXNERF_NETR_VERSION = 1 mutex kernel, 2x = round-1 v2
    // shapes, 3 (default) = the decoupled-role kernel of rays_roles.cuh
    static const int min_ctas = rtx_env_int("RTXNERF_NETR_CTAS", 2) == 3 ? 3 : 2;
    static const int version = rtx_env_int("RTXNERF_NETR_VERSION", 3) < 1 ? 3 : rtx_env_int("RTXNERF_NETR_VERSION", 3);
    if (version == 3) {
        RrArgs a = {};
        a.samples = d_samples;
        a.stride = sample_stride;
        a.out16 = d_out16;
        a.radiance4 = d_radiance4;
        a.encoded_out = d_encoded_out;
        a.num_hits = d_num_hits;
        a.indices = d_indices;
        a.n_rays = n_rays;
        a.S = samples_per_segment;
        a.out_activation = mlp->out_activation;
        return rr_launch(false, enc, mlp, lv, a, (cudaStream_t)stream);
    }
    {
        static RtxPerDeviceOnce a2, a3;
        RTX_CUDA_TRY((cudaError_t)rtx_set_max_smem(a2, network_forward_rays_kernel<2>, NetRSmem::BYTES));
        RTX_CUDA_TRY((cudaError_t)rtx_set_max_smem(a3, network_forward_rays_kernel<3>, NetRSmem::BYTES));
    }
    const long long groups = ((long long)n_rays + 31) / 32;
    if (version >= 2) {
#ifndef NETR2_BLOCKS_PER_SM
#define NETR2_BLOCKS_PER_SM 64
#endif
#define NETR2_LAUNCH(WGS, MINB)                                                                                     \
    do {                                                                                                            \
        static RtxPerDeviceOnce attr;                                                                               \
        RTX_CUDA_TRY((cudaError_t)rtx_set_max_smem(attr, network_forward_rays2_kernel<WGS, MINB>,                   \
                                                   NetR2Smem<WGS>::BYTES));                                         \
        long long nb = (groups + WGS - 1) / WGS;                                                                    \
        /* not a resident-only grid: one CTA per WGS groups (up to 64 per SM) lets the hardware scheduler       */ \
        /* balance the groups; the static grid-stride split over MINB CTAs per SM measured 5 % slower           */ \
        if (nb > (long long)RTX_NUM_SMS * NETR2_BLOCKS_PER_SM) nb = (long long)RTX_NUM_SMS * NETR2_BLOCKS_PER_SM;  \
        network_forward_rays2_kernel<WGS, MINB><<<(unsigned)nb, WGS * NET_TILE, NetR2Smem<WGS>::BYTES,              \
                                                  (cudaStream_t)stream>>>(                                          \
            lv, reinterpret_cast<const uint32_t*>(enc->table), mlp->weights, d_samples, sample_stride,              \
            mlp->out_activation, d_num_hits, d_indices, n_rays, samples_per_segment, d_out16, d_radiance4,          \
            d_encoded_out);                                                                                         \
    } while (0)
        if (version == 23) NETR2_LAUNCH(2, 3);        // 3 CTAs x 256 threads: 24 warps, 80 registers
        else if (version == 32) NETR2_LAUNCH(3, 2);   // 2 CTAs x 384 threads: 24 warps, 80 registers
        else if (version == 24) NETR2_LAUNCH(2, 4);   // 4 CTAs x 256 threads: 32 warps, 64 registers
        else NETR2_LAUNCH(4, 2);                      // 2 CTAs x 512 threads: 32 warps, 64 registers
#undef NETR2_LAUNCH
        return rtx_launch_status();
    }
    long long blocks = (groups + NETR_WGS - 1) / NETR_WGS;
    const long long resident = (long long)RTX_NUM_SMS * min_ctas;
    if (blocks > resident) blocks = resident;
    auto kern = min_ctas == 3 ? network_forward_rays_kernel<3> : network_forward_rays_kernel<2>;
    kern<<<(unsigned)blocks, NETR_THREADS, NetRSmem::BYTES, (cudaStream_t)stream>>>(
        lv, reinterpret_cast<const uint32_t*>(enc->table), mlp->weights, d_samples, sample_stride,
        mlp->out_activation, d_num_hits, d_indices, n_rays, samples_per_segment, d_out16, d_radiance4, d_encoded_out);
    return rtx_launch_status();
}

// The same forward without a sample buffer: sample (j,i) of ray x is rebuilt from the packed hit records of
// rtxnerf_ray_march_csr_seg / rtxnerf_batch_assemble_seg with the sampler's own arithmetic (sampler.cu:45-59,
// REGULAR sampling) — one 256-bit load per sample instead of three scalar loads from rows that lie a ray apart,
// and a training step needs no sampler launch.  Row order and results are those of rtxnerf_launch_sampler +
// rtxnerf_network_forward_rays.
extern "C" int rtxnerf_network_forward_seg(const rtxnerf_hashgrid* enc, const rtxnerf_mlp* mlp,
                                           const float* d_segments8, const int* d_num_hits, const int* d_indices,
                                           int n_rays, int samples_per_segment, uint16_t* d_out16,
                                           float* d_radiance4, uint16_t* d_encoded_out, void* stream)
{
    if (!enc || !mlp || !mlp->weights || !d_segments8 || !d_num_hits || !d_indices || (!d_out16 && !d_radiance4))
        return RTXNERF_EINVAL;
    if (n_rays < 0 || samples_per_segment < 1) return RTXNERF_EINVAL;
    if (((uintptr_t)d_segments8 & 31u) || ((uintptr_t)d_encoded_out & 31u)) return RTXNERF_EINVAL;
    if (mlp->in_dim != NET_IN || mlp->width != NET_W || mlp->n_hidden != 2) return RTXNERF_ECONFIG;
    if (mlp->out_activation != 0 && mlp->out_activation != 1) return RTXNERF_ECONFIG;
    Levels lv;
    int rc = levels_from(enc, &lv);
    if (rc) return rc;
    if (enc->n_levels * 2 > NET_IN) return RTXNERF_ECONFIG;
    if (n_rays == 0) return RTXNERF_OK;
    RrArgs a = {};
    a.seg8 = reinterpret_cast<const float4*>(d_segments8);
    a.out16 = d_out16;
    a.radiance4 = d_radiance4;
    a.encoded_out = d_encoded_out;
    a.num_hits = d_num_hits;
    a.indices = d_indices;
    a.n_rays = n_rays;
    a.S = samples_per_segment;
    a.out_activation = mlp->out_activation;
    return rr_launch(false, enc, mlp, lv, a, (cudaStream_t)stream);
}
