// network_backward.cu — backward pass of the hash-grid + fused MLP network for sm_100a.
//
// Replaces tiny-cuda-nn's network->backward as the reference calls it (main.cu:781) with the
// wiring the driver intends: dL/d(r,g,b,sigma) is the fp16 [N][4] buffer that
// launch_volrender_backward_cuda writes (vol_render.cu:136-139).  tiny-cuda-nn is absent from the
// reference archive: the algorithm is restated (PARITY UNPINNED) —
//   dOut_j   = g_j * act'(out_j)            (sigmoid: s(1-s), recomputed from the forward)
//   dH_k     = (dH_{k+1} . W_{k+1}) masked by ReLU'(H_k)
//   dW_k    += dH_{k+1}^T . H_k             (summed over all samples)
//   dEnc     = dH_1 . W_0                    -> scattered into the hash table by the trilinear weights
//
// Two kernels:
//  (1) mlp_backward_kernel: one CTA = one 128-sample tile at a time on tcgen05.  The forward is
//      recomputed from the encoded features the training forward stored (64 B/sample) instead of
//      saving 256 B/sample of hidden activations.  Every tile in shared memory is read by the
//      tensor cores in both majors (tcgen05.cuh): [sample][feature] is the K-major A of the
//      dgrad GEMM and the MN-major operand of the weight-gradient GEMM, the resident weight tile
//      [out][in] is K-major for the forward and MN-major for its dgrad — no transposed copies.
//      Weight gradients accumulate in TMEM across ALL tiles of the CTA (M = 64 accumulators) and
//      leave once, as fp32 atomics, at the end.
//  (2) hashgrid_backward_kernel: ray-coherent scatter (lane = ray, warp = sample slot, as in
//      network_rays.cu) of dEnc into the fp32 table gradient with 8-byte vector reductions.
#include <math.h>
#include <stdlib.h>

#include "common.cuh"
#include "hashgrid.cuh"
#include "mlp_tile.cuh"

#define NB_TMEM_COLS 256
#define NB_D 0        // [128 x 64] forward / dgrad accumulator
#define NB_G1 64      // dW1   [64 out x 64 in]
#define NB_G0 128     // dW0   [64 out x 32 in]
#define NB_G2 160     // dW2^T [64 in  x 16 out]
// A operands of the forward / dgrad products (TS form, tcgen05.cuh mma_f16_ts): the tile that is the A of the next
// product is ALSO written to TMEM by the epilogue that produces it.  An MMA with both operands in shared memory
// costs (A + B bytes) / 128 cycles of the SM's shared-memory port (scripts/probes/mma_issue_probe.cu) and this
// kernel's epilogues need the same port; with A in TMEM the six products read only their weight tile there:
// 1 488 -> 880 cycles of operand reads per 128-sample tile.  The shared-memory copies stay: they are the
// MN-major operands of the weight-gradient products.
#define NB_AH 176     // [128 x 64] fp16: H1, H2, dH2, dH1 in turn (32 columns)
#define NB_A0 208     // [128 x 32] fp16 encoded features (16 columns)
#define NB_ADO 224    // [128 x 16] fp16 dOut (8 columns)

struct NbSmem {
    static constexpr int W0 = 0;                             // [64][32]
    static constexpr int W1 = W0 + NET_W * NET_IN * 2;       // [64][64]
    static constexpr int W2 = W1 + NET_W * NET_W * 2;        // [16][64]
    static constexpr int A0 = W2 + NET_OUT * NET_W * 2;      // [128][32] encoded features
    static constexpr int H1 = A0 + NET_TILE * NET_IN * 2;    // [128][64]
    static constexpr int H2 = H1 + NET_TILE * NET_W * 2;     // [128][64], later dH1
    static constexpr int DO = H2 + NET_TILE * NET_W * 2;     // [128][16] dOut
    static constexpr int DH = DO + NET_TILE * NET_OUT * 2;   // [128][64] dH2
    static constexpr int BAR = DH + NET_TILE * NET_W * 2;
    static constexpr int BYTES = BAR + 16;
};

// K-major A [128 x K] (tile of C = K columns) times K-major B [N x K]: the forward layers
template <int N, int K>
__device__ __forceinline__ void mma_fwd(uint32_t d, uint32_t a, uint32_t w)
{
    constexpr uint32_t id = tc::idesc_f16_f32_major(NET_TILE, N, 0, 0);
    const uint64_t da = tc::smem_desc(a, 128, K * 16), dw = tc::smem_desc(w, 128, K * 16);
#pragma unroll
    for (int ks = 0; ks < K / 16; ks++)
        tc::mma_f16_ss(d, tc::smem_desc_advance(da, ks * 256), tc::smem_desc_advance(dw, ks * 256), id, ks > 0 ? 1u : 0u);
}
// dgrad: D[128 x N_in] = dH[128 x K_out] (K-major, tile of K_out columns) . W[K_out][N_in] (MN-major B)
template <int N_IN, int K_OUT>
__device__ __forceinline__ void mma_dgrad(uint32_t d, uint32_t dh, uint32_t w)
{
    constexpr uint32_t id = tc::idesc_f16_f32_major(NET_TILE, N_IN, 0, 1);
    const uint64_t dd = tc::smem_desc(dh, 128, K_OUT * 16), dw = tc::smem_desc(w, N_IN * 16, 128);
#pragma unroll
    for (int ks = 0; ks < K_OUT / 16; ks++)
        tc::mma_f16_ss(d, tc::smem_desc_advance(dd, ks * 256), tc::smem_desc_advance(dw, ks * 2 * (N_IN * 16)), id,
                       ks > 0 ? 1u : 0u);
}
// wgrad: G[64 x N] (+)= X[128 samples][64]^T . Y[128 samples][N]   (both MN-major, K = samples)
template <int N>
__device__ __forceinline__ void mma_wgrad(uint32_t g, uint32_t x, uint32_t y, bool accumulate)
{
    constexpr uint32_t id = tc::idesc_f16_f32_major(64, N, 1, 1);
    const uint64_t dx = tc::smem_desc(x, NET_W * 16, 128), dy = tc::smem_desc(y, N * 16, 128);
#pragma unroll
    for (int ks = 0; ks < NET_TILE / 16; ks++)
        tc::mma_f16_ss(g, tc::smem_desc_advance(dx, ks * 2 * (NET_W * 16)), tc::smem_desc_advance(dy, ks * 2 * (N * 16)), id,
                       (accumulate || ks > 0) ? 1u : 0u);
}

// TS forms: A [128 x K] from TMEM (K / 2 columns), B from shared memory
template <int N, int K>
__device__ __forceinline__ void mma_fwd_ts(uint32_t d, uint32_t a_tmem, uint32_t w)
{
    constexpr uint32_t id = tc::idesc_f16_f32_major(NET_TILE, N, 0, 0);
    const uint64_t dw = tc::smem_desc(w, 128, K * 16);
#pragma unroll
    for (int ks = 0; ks < K / 16; ks++)
        tc::mma_f16_ts(d, a_tmem + ks * 8, tc::smem_desc_advance(dw, ks * 256), id, ks > 0 ? 1u : 0u);
}
template <int N_IN, int K_OUT>
__device__ __forceinline__ void mma_dgrad_ts(uint32_t d, uint32_t a_tmem, uint32_t w)
{
    constexpr uint32_t id = tc::idesc_f16_f32_major(NET_TILE, N_IN, 0, 1);
    const uint64_t dw = tc::smem_desc(w, N_IN * 16, 128);
#pragma unroll
    for (int ks = 0; ks < K_OUT / 16; ks++)
        tc::mma_f16_ts(d, a_tmem + ks * 8, tc::smem_desc_advance(dw, ks * 2 * (N_IN * 16)), id, ks > 0 ? 1u : 0u);
}

// forward epilogue of this kernel: relu_row_to_smem (mlp_tile.cuh) plus the TMEM copy of the row (the next A)
__device__ __forceinline__ void relu_row_to_smem_tmem(uint32_t taddr, uint8_t* a_next, int row, uint32_t t_a)
{
#pragma unroll
    for (int h = 0; h < NET_W / 32; h++) {
        uint32_t r0[16], r1[16];
        tc::tmem_ld_x16(taddr + h * 32, r0);
        tc::tmem_ld_x16(taddr + h * 32 + 16, r1);
        tc::tmem_ld_wait();
        uint32_t pk[16];
#pragma unroll
        for (int j = 0; j < 8; j++) {
            pk[j] = relu_pack_f16x2(__uint_as_float(r0[2 * j]), __uint_as_float(r0[2 * j + 1]));
            pk[8 + j] = relu_pack_f16x2(__uint_as_float(r1[2 * j]), __uint_as_float(r1[2 * j + 1]));
        }
        tc::tmem_st_x16(t_a + h * 16, pk);
#pragma unroll
        for (int c = 0; c < 4; c++)
            *reinterpret_cast<uint4*>(a_next + tile_chunk_off(row, 4 * h + c, NET_W)) =
                make_uint4(pk[4 * c], pk[4 * c + 1], pk[4 * c + 2], pk[4 * c + 3]);
    }
}

// dgrad epilogue: TMEM row -> masked by ReLU'(h) (h = this row of the forward tile) -> fp16 tile
__device__ __forceinline__ void masked_row_to_smem(uint32_t taddr, const uint8_t* h_tile, uint8_t* dst, int row,
                                                   uint32_t t_a)
{
#pragma unroll
    for (int q = 0; q < NET_W / 16; q++) {
        uint32_t r[16];
        tc::tmem_ld_x16(taddr + q * 16, r);
        tc::tmem_ld_wait();
        const uint4 h0 = *reinterpret_cast<const uint4*>(h_tile + tile_chunk_off(row, 2 * q, NET_W));
        const uint4 h1 = *reinterpret_cast<const uint4*>(h_tile + tile_chunk_off(row, 2 * q + 1, NET_W));
        const uint32_t hw[8] = {h0.x, h0.y, h0.z, h0.w, h1.x, h1.y, h1.z, h1.w};
        uint32_t pk[8];
#pragma unroll
        for (int j = 0; j < 8; j++) {
            // H is a ReLU output: "> 0" is "nonzero" on the fp16 bits (no negatives, -0 never stored)
            const float a = (hw[j] & 0xffffu) ? __uint_as_float(r[2 * j]) : 0.0f;
            const float b = (hw[j] >> 16) ? __uint_as_float(r[2 * j + 1]) : 0.0f;
            const __half2 v = __floats2half2_rn(a, b);
            pk[j] = *reinterpret_cast<const uint32_t*>(&v);
        }
        tc::tmem_st_x8(t_a + q * 8, pk);
        *reinterpret_cast<uint4*>(dst + tile_chunk_off(row, 2 * q, NET_W)) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
        *reinterpret_cast<uint4*>(dst + tile_chunk_off(row, 2 * q + 1, NET_W)) = make_uint4(pk[4], pk[5], pk[6], pk[7]);
    }
}

__global__ void __launch_bounds__(NET_TILE)
mlp_backward_kernel(const uint16_t* __restrict__ weights, const uint16_t* __restrict__ encoded,
                    const uint16_t* __restrict__ dl_dradiance, int out_activation,
                    const long long* __restrict__ d_count, long long count_scale, long long n_max,
                    uint16_t* __restrict__ d_encoded, float* __restrict__ grad_weights)
{
    extern __shared__ __align__(128) uint8_t smem[];
    uint64_t* bar = reinterpret_cast<uint64_t*>(smem + NbSmem::BAR);
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + NbSmem::BAR + 8);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;

    load_weight_tile(smem + NbSmem::W0, weights, NET_W, NET_IN, tid, NET_TILE);
    load_weight_tile(smem + NbSmem::W1, weights + NET_W * NET_IN, NET_W, NET_W, tid, NET_TILE);
    load_weight_tile(smem + NbSmem::W2, weights + NET_W * NET_IN + NET_W * NET_W, NET_OUT, NET_W, tid, NET_TILE);
    if (tid == 0) {
        tc::mbar_init(bar, 1);
        tc::fence_mbar_init();
    }
    if (warp == 0) tc::tmem_alloc<NB_TMEM_COLS>(tmem_slot);
    tc::fence_proxy_async_smem();
    tc::fence_before_thread_sync();
    __syncthreads();
    tc::fence_after_thread_sync();
    const uint32_t tm = *tmem_slot;
    const uint32_t t_row = tm + ((uint32_t)(warp * 32) << 16);
    const uint32_t s_w0 = tc::smem_u32(smem + NbSmem::W0), s_w1 = tc::smem_u32(smem + NbSmem::W1),
                   s_w2 = tc::smem_u32(smem + NbSmem::W2), s_a0 = tc::smem_u32(smem + NbSmem::A0),
                   s_h1 = tc::smem_u32(smem + NbSmem::H1), s_h2 = tc::smem_u32(smem + NbSmem::H2),
                   s_do = tc::smem_u32(smem + NbSmem::DO), s_dh = tc::smem_u32(smem + NbSmem::DH);
    uint32_t phase = 0;
    bool have_grads = false;   // the TMEM weight-gradient accumulators hold at least one tile

    auto sync_then = [&](auto issue) {   // publish this phase's smem tile + TMEM operand, run the MMAs, wait for them
        tc::tmem_st_wait();
        tc::fence_proxy_async_smem();
        tc::fence_before_thread_sync();
        __syncthreads();
        if (warp == 0 && tc::elect_one()) {     // one lane of the converged warp 0 (tcgen05.cuh: elect_one)
            tc::fence_after_thread_sync();
            issue();
            tc::mma_commit(bar);
        }
        tc::mbar_wait(bar, phase);
        phase ^= 1;
        tc::fence_after_thread_sync();
    };

    const long long n = row_count(d_count, count_scale, n_max);
    const long long n_tiles = (n + NET_TILE - 1) / NET_TILE;
    uint4 pre[NET_IN / 8];   // software pipeline: the next tile's inputs are in flight during this tile's layers
    uint2 pre_g;
    auto fetch_row = [&](long long t) {
        const long long sn = t * NET_TILE + tid;
        const bool ok = sn < n;
        const uint4* src = reinterpret_cast<const uint4*>(encoded + sn * NET_IN);
#pragma unroll
        for (int c = 0; c < NET_IN / 8; c++) pre[c] = ok ? __ldg(src + c) : make_uint4(0, 0, 0, 0);
        pre_g = ok ? __ldg(reinterpret_cast<const uint2*>(dl_dradiance + sn * 4)) : make_uint2(0, 0);
    };
    fetch_row(blockIdx.x);
    for (long long tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        const long long s = tile * NET_TILE + tid;
        const bool valid = s < n;
        {   // A0: the stored encoded features of this row (fetched during the previous tile)
#pragma unroll
            for (int c = 0; c < NET_IN / 8; c++)
                *reinterpret_cast<uint4*>(smem + NbSmem::A0 + tile_chunk_off(tid, c, NET_IN)) = pre[c];
            const uint32_t a0w[16] = {pre[0].x, pre[0].y, pre[0].z, pre[0].w, pre[1].x, pre[1].y, pre[1].z, pre[1].w,
                                      pre[2].x, pre[2].y, pre[2].z, pre[2].w, pre[3].x, pre[3].y, pre[3].z, pre[3].w};
            tc::tmem_st_x16(t_row + NB_A0, a0w);
        }
        const uint2 gb = pre_g;
        fetch_row(tile + gridDim.x);
        // ---- forward recompute
        sync_then([&] { mma_fwd_ts<NET_W, NET_IN>(tm + NB_D, tm + NB_A0, s_w0); });
        relu_row_to_smem_tmem(t_row + NB_D, smem + NbSmem::H1, tid, t_row + NB_AH);
        sync_then([&] { mma_fwd_ts<NET_W, NET_W>(tm + NB_D, tm + NB_AH, s_w1); });
        relu_row_to_smem_tmem(t_row + NB_D, smem + NbSmem::H2, tid, t_row + NB_AH);
        sync_then([&] { mma_fwd_ts<NET_OUT, NET_W>(tm + NB_D, tm + NB_AH, s_w2); });
        {   // dOut = g * act'(out) on (r,g,b,sigma); the 12 padding outputs carry no gradient
            uint32_t r[16];
            output_row_load(t_row + NB_D, r);
            float d[4] = {0.f, 0.f, 0.f, 0.f};
            if (valid) {
                const float2 g01 = h2_bits_to_f2(gb.x), g23 = h2_bits_to_f2(gb.y);
                const float g[4] = {g01.x, g01.y, g23.x, g23.y};
#pragma unroll
                for (int j = 0; j < 4; j++) {
                    float da = 1.0f;
                    if (out_activation == 1) {
                        const float sg = __fdividef(1.0f, 1.0f + __expf(-__uint_as_float(r[j])));
                        da = sg * (1.0f - sg);
                    }
                    d[j] = g[j] * da;
                }
            }
            const __half2 d01 = __floats2half2_rn(d[0], d[1]), d23 = __floats2half2_rn(d[2], d[3]);
            *reinterpret_cast<uint4*>(smem + NbSmem::DO + tile_chunk_off(tid, 0, NET_OUT)) =
                make_uint4(*reinterpret_cast<const uint32_t*>(&d01), *reinterpret_cast<const uint32_t*>(&d23), 0u, 0u);
            *reinterpret_cast<uint4*>(smem + NbSmem::DO + tile_chunk_off(tid, 1, NET_OUT)) = make_uint4(0u, 0u, 0u, 0u);
            const uint32_t dow[8] = {*reinterpret_cast<const uint32_t*>(&d01), *reinterpret_cast<const uint32_t*>(&d23),
                                     0u, 0u, 0u, 0u, 0u, 0u};
            tc::tmem_st_x8(t_row + NB_ADO, dow);
        }
        // ---- layer 2: dW2^T += H2^T . dOut ;  dH2 = (dOut . W2) * ReLU'(H2)
        sync_then([&] {
            mma_wgrad<NET_OUT>(tm + NB_G2, s_h2, s_do, have_grads);
            mma_dgrad_ts<NET_W, NET_OUT>(tm + NB_D, tm + NB_ADO, s_w2);
        });
        masked_row_to_smem(t_row + NB_D, smem + NbSmem::H2, smem + NbSmem::DH, tid, t_row + NB_AH);
        // ---- layer 1: dW1 += dH2^T . H1 ;  dH1 = (dH2 . W1) * ReLU'(H1)   (dH1 overwrites H2)
        sync_then([&] {
            mma_wgrad<NET_W>(tm + NB_G1, s_dh, s_h1, have_grads);
            mma_dgrad_ts<NET_W, NET_W>(tm + NB_D, tm + NB_AH, s_w1);
        });
        masked_row_to_smem(t_row + NB_D, smem + NbSmem::H1, smem + NbSmem::H2, tid, t_row + NB_AH);
        // ---- layer 0: dW0 += dH1^T . A0 ;  dEnc = dH1 . W0
        sync_then([&] {
            mma_wgrad<NET_IN>(tm + NB_G0, s_h2, s_a0, have_grads);
            mma_dgrad_ts<NET_IN, NET_W>(tm + NB_D, tm + NB_AH, s_w0);
        });
        have_grads = true;
#pragma unroll
        for (int q = 0; q < NET_IN / 16; q++) {
            uint32_t r[16];
            tc::tmem_ld_x16(t_row + NB_D + q * 16, r);
            tc::tmem_ld_wait();
            if (valid) {
                uint32_t pk[8];
#pragma unroll
                for (int j = 0; j < 8; j++) {
                    const __half2 v = __floats2half2_rn(__uint_as_float(r[2 * j]), __uint_as_float(r[2 * j + 1]));
                    pk[j] = *reinterpret_cast<const uint32_t*>(&v);
                }
                uint4* o = reinterpret_cast<uint4*>(d_encoded + s * NET_IN + q * 16);
                o[0] = make_uint4(pk[0], pk[1], pk[2], pk[3]);
                o[1] = make_uint4(pk[4], pk[5], pk[6], pk[7]);
            }
        }
    }
    // ---- weight gradients leave TMEM once.  An M = 64 accumulator keeps row r in TMEM lane
    // (r / 16) * 32 + r % 16: lanes 0..15 of each warp's quarter (umma_probe.cu).
    tc::fence_before_thread_sync();
    __syncthreads();
    tc::fence_after_thread_sync();
    if (have_grads) {
        const int row = warp * 16 + lane;   // meaningful for lane < 16
        float* g0 = grad_weights;
        float* g1 = g0 + NET_W * NET_IN;
        float* g2 = g1 + NET_W * NET_W;
#pragma unroll
        for (int q = 0; q < NET_W / 16; q++) {
            uint32_t r[16];
            tc::tmem_ld_x16(t_row + NB_G1 + q * 16, r);
            tc::tmem_ld_wait();
            if (lane < 16)
                for (int j = 0; j < 16; j++) atomicAdd(g1 + row * NET_W + q * 16 + j, __uint_as_float(r[j]));
        }
#pragma unroll
        for (int q = 0; q < NET_IN / 16; q++) {
            uint32_t r[16];
            tc::tmem_ld_x16(t_row + NB_G0 + q * 16, r);
            tc::tmem_ld_wait();
            if (lane < 16)
                for (int j = 0; j < 16; j++) atomicAdd(g0 + row * NET_IN + q * 16 + j, __uint_as_float(r[j]));
        }
        {
            uint32_t r[16];
            tc::tmem_ld_x16(t_row + NB_G2, r);
            tc::tmem_ld_wait();
            if (lane < 16)
                for (int j = 0; j < 16; j++) atomicAdd(g2 + j * NET_W + row, __uint_as_float(r[j]));   // [out j][in row]
        }
    }
    tc::fence_before_thread_sync();
    __syncthreads();
    tc::fence_after_thread_sync();
    if (warp == 0) tc::tmem_dealloc<NB_TMEM_COLS>(tm);
}

// ---------------------------------------------------------------------------------------
// (1b) mlp_backward with TWO tiles per CTA (two CTAs per SM: four tiles and four issuing threads per SM).  The
// kernel above is bound by the latency of its serial phase chain (barrier -> issue -> MMA -> commit -> wait ->
// epilogue, ~1 000 cycles x 6 per tile) with the MMA pipe ~40 % busy; what limits the tiles in flight is TMEM: the
// weight-gradient accumulators take 112 of a CTA's 256 columns.  Here two warp-groups of a CTA walk their own
// tiles through the same six phases, each with its own accumulator D, its own issuing thread and its own named
// barrier, and SHARE the three weight-gradient accumulators: MMAs issued by different threads into one TMEM
// accumulator are serialised by the tensor pipe without losing updates (scripts/probes/mma_shared_acc_probe.cu:
// 1 / 2 / 4 issuers, 16 384 accumulating MMAs, exact sums).  The accumulators are zeroed once with tcgen05.st, so
// every weight-gradient MMA accumulates.  Shared memory per tile drops to 44 KB (dH2 overwrites H2 and dH1
// overwrites H1 in place: a thread rewrites the chunks of its own row after the MMAs that read them have completed),
// 102 KB per CTA.  Both operands of every product come from shared memory (the TS form of the kernel above would
// need 2 x 56 more TMEM columns).
// ---------------------------------------------------------------------------------------
#define N2_GROUPS 2
#define N2_THREADS (N2_GROUPS * NET_TILE)
#define N2_TMEM_COLS 256
#define N2_D(g) (64 * (g))
#define N2_G1 128
#define N2_G0 192
#define N2_G2 224
struct N2Smem {
    static constexpr int W0 = 0;
    static constexpr int W1 = W0 + NET_W * NET_IN * 2;
    static constexpr int W2 = W1 + NET_W * NET_W * 2;
    static constexpr int SLOT = W2 + NET_OUT * NET_W * 2;
    static constexpr int A0 = 0;                               // inside a slot: [128][32] encoded features
    static constexpr int H1 = A0 + NET_TILE * NET_IN * 2;      // [128][64], later dH1
    static constexpr int H2 = H1 + NET_TILE * NET_W * 2;       // [128][64], later dH2
    static constexpr int DO = H2 + NET_TILE * NET_W * 2;       // [128][16] dOut
    static constexpr int SLOT_BYTES = DO + NET_TILE * NET_OUT * 2;
    static constexpr int BAR = SLOT + N2_GROUPS * SLOT_BYTES;
    static constexpr int BYTES = BAR + N2_GROUPS * 8 + 16;
};

// dgrad epilogue in place: TMEM row -> masked by ReLU'(h), h = this row of `tile` -> fp16 over the same chunks
__device__ __forceinline__ void masked_row_in_place(uint32_t taddr, uint8_t* tile, int row)
{
#pragma unroll
    for (int q = 0; q < NET_W / 16; q++) {
        uint32_t r[16];
        tc::tmem_ld_x16(taddr + q * 16, r);
        uint4* c0 = reinterpret_cast<uint4*>(tile + tile_chunk_off(row, 2 * q, NET_W));
        uint4* c1 = reinterpret_cast<uint4*>(tile + tile_chunk_off(row, 2 * q + 1, NET_W));
        const uint4 h0 = *c0, h1 = *c1;
        tc::tmem_ld_wait();
        const uint32_t hw[8] = {h0.x, h0.y, h0.z, h0.w, h1.x, h1.y, h1.z, h1.w};
        uint32_t pk[8];
#pragma unroll
        for (int j = 0; j < 8; j++) {
            const float a = (hw[j] & 0xffffu) ? __uint_as_float(r[2 * j]) : 0.0f;
            const float b = (hw[j] >> 16) ? __uint_as_float(r[2 * j + 1]) : 0.0f;
            const __half2 v = __floats2half2_rn(a, b);
            pk[j] = *reinterpret_cast<const uint32_t*>(&v);
        }
        *c0 = make_uint4(pk[0], pk[1], pk[2], pk[3]);
        *c1 = make_uint4(pk[4], pk[5], pk[6], pk[7]);
    }
}

__global__ void __launch_bounds__(N2_THREADS, 2)
mlp_backward2_kernel(const uint16_t* __restrict__ weights, const uint16_t* __restrict__ encoded,
                     const uint16_t* __restrict__ dl_dradiance, int out_activation,
                     const long long* __restrict__ d_count, long long count_scale, long long n_max,
                     uint16_t* __restrict__ d_encoded, float* __restrict__ grad_weights)
{
    extern __shared__ __align__(128) uint8_t smem[];
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + N2Smem::BAR);
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + N2_GROUPS);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int grp = warp >> 2, row = tid & (NET_TILE - 1);

    load_weight_tile(smem + N2Smem::W0, weights, NET_W, NET_IN, tid, N2_THREADS);
    load_weight_tile(smem + N2Smem::W1, weights + NET_W * NET_IN, NET_W, NET_W, tid, N2_THREADS);
    load_weight_tile(smem + N2Smem::W2, weights + NET_W * NET_IN + NET_W * NET_W, NET_OUT, NET_W, tid, N2_THREADS);
    if (tid == 0) {
        for (int g = 0; g < N2_GROUPS; g++) tc::mbar_init(bars + g, 1);
        tc::fence_mbar_init();
    }
    if (warp == 0) tc::tmem_alloc<N2_TMEM_COLS>(tmem_slot);
    tc::fence_proxy_async_smem();
    tc::fence_before_thread_sync();
    __syncthreads();
    tc::fence_after_thread_sync();
    const uint32_t tm = *tmem_slot;
    const uint32_t t_q = tm + ((uint32_t)((warp & 3) * 32) << 16);      // this warp's TMEM lane quarter
    if (grp == 0) {      // the shared weight-gradient accumulators start at zero: every wgrad MMA accumulates
        uint32_t z[16];
#pragma unroll
        for (int i = 0; i < 16; i++) z[i] = 0u;
#pragma unroll
        for (int c = N2_G1; c < N2_G2 + NET_OUT; c += 16) tc::tmem_st_x16(t_q + c, z);
        tc::tmem_st_wait();
    }
    tc::fence_before_thread_sync();
    __syncthreads();
    tc::fence_after_thread_sync();

    uint8_t* sl = smem + N2Smem::SLOT + grp * N2Smem::SLOT_BYTES;
    const uint32_t t_d = t_q + N2_D(grp);
    const uint32_t d_col = tm + N2_D(grp);
    const uint32_t s_w0 = tc::smem_u32(smem + N2Smem::W0), s_w1 = tc::smem_u32(smem + N2Smem::W1),
                   s_w2 = tc::smem_u32(smem + N2Smem::W2), s_a0 = tc::smem_u32(sl + N2Smem::A0),
                   s_h1 = tc::smem_u32(sl + N2Smem::H1), s_h2 = tc::smem_u32(sl + N2Smem::H2),
                   s_do = tc::smem_u32(sl + N2Smem::DO);
    uint64_t* bar = bars + grp;
    uint32_t phase = 0;
    const bool issuer_warp = (warp & 3) == 0;

    auto sync_then = [&](auto issue) {   // publish the group's tiles, run its MMAs, wait for them
        tc::fence_proxy_async_smem();
        tc::fence_before_thread_sync();
        asm volatile("bar.sync %0, %1;" ::"r"(1 + grp), "n"(NET_TILE) : "memory");
        if (issuer_warp && tc::elect_one()) {
            tc::fence_after_thread_sync();
            issue();
            tc::mma_commit(bar);
        }
        tc::mbar_wait(bar, phase);
        phase ^= 1;
        tc::fence_after_thread_sync();
    };

    const long long n = row_count(d_count, count_scale, n_max);
    const long long n_tiles = (n + NET_TILE - 1) / NET_TILE;
    const long long stride = (long long)gridDim.x * N2_GROUPS;
    const long long first = (long long)blockIdx.x * N2_GROUPS + grp;
    uint4 pre[NET_IN / 8];
    uint2 pre_g;
    auto fetch_row = [&](long long t) {
        const long long sn = t * NET_TILE + row;
        const bool ok = sn < n;
        const uint4* src = reinterpret_cast<const uint4*>(encoded + sn * NET_IN);
#pragma unroll
        for (int c = 0; c < NET_IN / 8; c++) pre[c] = ok ? __ldg(src + c) : make_uint4(0, 0, 0, 0);
        pre_g = ok ? __ldg(reinterpret_cast<const uint2*>(dl_dradiance + sn * 4)) : make_uint2(0, 0);
    };
    fetch_row(first);
    for (long long tile = first; tile < n_tiles; tile += stride) {
        const long long s = tile * NET_TILE + row;
        const bool valid = s < n;
#pragma unroll
        for (int c = 0; c < NET_IN / 8; c++)
            *reinterpret_cast<uint4*>(sl + N2Smem::A0 + tile_chunk_off(row, c, NET_IN)) = pre[c];
        const uint2 gb = pre_g;
        fetch_row(tile + stride);
        // ---- forward recompute
        sync_then([&] { mma_fwd<NET_W, NET_IN>(d_col, s_a0, s_w0); });
        relu_row_to_smem(t_d, sl + N2Smem::H1, row);
        sync_then([&] { mma_fwd<NET_W, NET_W>(d_col, s_h1, s_w1); });
        relu_row_to_smem(t_d, sl + N2Smem::H2, row);
        sync_then([&] { mma_fwd<NET_OUT, NET_W>(d_col, s_h2, s_w2); });
        {   // dOut = g * act'(out) on (r,g,b,sigma); the 12 padding outputs carry no gradient
            uint32_t r[16];
            output_row_load(t_d, r);
            float d[4] = {0.f, 0.f, 0.f, 0.f};
            if (valid) {
                const float2 g01 = h2_bits_to_f2(gb.x), g23 = h2_bits_to_f2(gb.y);
                const float g[4] = {g01.x, g01.y, g23.x, g23.y};
#pragma unroll
                for (int j = 0; j < 4; j++) {
                    float da = 1.0f;
                    if (out_activation == 1) {
                        const float sg = __fdividef(1.0f, 1.0f + __expf(-__uint_as_float(r[j])));
                        da = sg * (1.0f - sg);
                    }
                    d[j] = g[j] * da;
                }
            }
            const __half2 d01 = __floats2half2_rn(d[0], d[1]), d23 = __floats2half2_rn(d[2], d[3]);
            *reinterpret_cast<uint4*>(sl + N2Smem::DO + tile_chunk_off(row, 0, NET_OUT)) =
                make_uint4(*reinterpret_cast<const uint32_t*>(&d01), *reinterpret_cast<const uint32_t*>(&d23), 0u, 0u);
            *reinterpret_cast<uint4*>(sl + N2Smem::DO + tile_chunk_off(row, 1, NET_OUT)) = make_uint4(0u, 0u, 0u, 0u);
        }
        // ---- layer 2: dW2^T += H2^T . dOut ;  dH2 = (dOut . W2) * ReLU'(H2), over H2
        sync_then([&] {
            mma_wgrad<NET_OUT>(tm + N2_G2, s_h2, s_do, true);
            mma_dgrad<NET_W, NET_OUT>(d_col, s_do, s_w2);
        });
        masked_row_in_place(t_d, sl + N2Smem::H2, row);
        // ---- layer 1: dW1 += dH2^T . H1 ;  dH1 = (dH2 . W1) * ReLU'(H1), over H1
        sync_then([&] {
            mma_wgrad<NET_W>(tm + N2_G1, s_h2, s_h1, true);
            mma_dgrad<NET_W, NET_W>(d_col, s_h2, s_w1);
        });
        masked_row_in_place(t_d, sl + N2Smem::H1, row);
        // ---- layer 0: dW0 += dH1^T . A0 ;  dEnc = dH1 . W0
        sync_then([&] {
            mma_wgrad<NET_IN>(tm + N2_G0, s_h1, s_a0, true);
            mma_dgrad<NET_IN, NET_W>(d_col, s_h1, s_w0);
        });
#pragma unroll
        for (int q = 0; q < NET_IN / 16; q++) {
            uint32_t r[16];
            tc::tmem_ld_x16(t_d + q * 16, r);
            tc::tmem_ld_wait();
            if (valid) {
                uint32_t pk[8];
#pragma unroll
                for (int j = 0; j < 8; j++) {
                    const __half2 v = __floats2half2_rn(__uint_as_float(r[2 * j]), __uint_as_float(r[2 * j + 1]));
                    pk[j] = *reinterpret_cast<const uint32_t*>(&v);
                }
                uint4* o = reinterpret_cast<uint4*>(d_encoded + s * NET_IN + q * 16);
                o[0] = make_uint4(pk[0], pk[1], pk[2], pk[3]);
                o[1] = make_uint4(pk[4], pk[5], pk[6], pk[7]);
            }
        }
    }
    // ---- weight gradients leave TMEM once (M = 64 accumulator: row r in TMEM lane (r / 16) * 32 + r % 16)
    tc::fence_before_thread_sync();
    __syncthreads();
    tc::fence_after_thread_sync();
    if (grp == 0 && (long long)blockIdx.x * N2_GROUPS < n_tiles) {
        const int grow = warp * 16 + lane;   // meaningful for lane < 16
        float* g0 = grad_weights;
        float* g1 = g0 + NET_W * NET_IN;
        float* g2 = g1 + NET_W * NET_W;
#pragma unroll
        for (int q = 0; q < NET_W / 16; q++) {
            uint32_t r[16];
            tc::tmem_ld_x16(t_q + N2_G1 + q * 16, r);
            tc::tmem_ld_wait();
            if (lane < 16)
                for (int j = 0; j < 16; j++) atomicAdd(g1 + grow * NET_W + q * 16 + j, __uint_as_float(r[j]));
        }
#pragma unroll
        for (int q = 0; q < NET_IN / 16; q++) {
            uint32_t r[16];
            tc::tmem_ld_x16(t_q + N2_G0 + q * 16, r);
            tc::tmem_ld_wait();
            if (lane < 16)
                for (int j = 0; j < 16; j++) atomicAdd(g0 + grow * NET_IN + q * 16 + j, __uint_as_float(r[j]));
        }
        {
            uint32_t r[16];
            tc::tmem_ld_x16(t_q + N2_G2, r);
            tc::tmem_ld_wait();
            if (lane < 16)
                for (int j = 0; j < 16; j++) atomicAdd(g2 + j * NET_W + grow, __uint_as_float(r[j]));   // [out j][in row]
        }
    }
    tc::fence_before_thread_sync();
    __syncthreads();
    tc::fence_after_thread_sync();
    if (warp == 0) tc::tmem_dealloc<N2_TMEM_COLS>(tm);
}

// ---------------------------------------------------------------------------------------
// (2) dEnc -> table gradient.  grad_table is fp32 [n_entries][2]; a corner update is one 8-byte
// vector reduction, an x-pair of corners that shares an aligned entry pair one 16-byte one.
// lane = ray, warp = sample slot (as in network_rays.cu): the 32 lanes of a warp are
// neighbouring rays at the same depth, so consecutive lanes sit in the SAME cell — all 32 on the
// coarse levels, runs of 2-8 on the fine ones — and would send their reductions to the same
// addresses, which L1TEX/L2 serialise (first version: 94 ms for 2^18 rays x 128 samples; ncu:
// L1TEX 89 % busy).  Per level the warp therefore sums the 16 corner contributions over each
// run of consecutive lanes with equal cell (a segmented shuffle reduction whose run structure
// comes from one ballot; it stops at the longest run) and only the head lane of a run issues
// the reductions.
// Tried and dropped (round 2): cutting long runs into pieces of at most 8 / 16 lanes to bound the shuffle rounds
// (56 shuffles per level and warp on average, MIO the first stall reason).  It moves the load from the shuffles
// to the L2 reductions, which are as close to their limit: 10.17 -> 9.99 ms on one training view, 15.3 -> 17.2 ms
// on the store-fed batches of the bench (profiles/r02_bb_scatter_run_cap.txt).  Timed alone (a level mask), the
// two coarsest levels cost 1.5 and 1.6 ms of the 5.8 ms and a fine level 0.4 ms: reductions on the same few
// entries serialise in L2 (profiles/r02_bc_scatter_per_level.txt).  Collecting the coarse levels in a
// shared-memory accumulator of the CTA (open addressing, shared-memory atomics, one flush per ray group) moved
// that serialisation into shared memory and lost 2.5 ms (profiles/r02_bd_scatter_cta_accumulator.txt).
// ---------------------------------------------------------------------------------------
#define HB_WARPS 8
#ifndef HB_MIN_BLOCKS
#define HB_MIN_BLOCKS 4
#endif
__global__ void __launch_bounds__(HB_WARPS * 32, HB_MIN_BLOCKS)
hashgrid_backward_kernel(Levels lv, const float* __restrict__ samples, int stride, const float* __restrict__ seg8,
                         const uint16_t* __restrict__ d_encoded, const int* __restrict__ num_hits,
                         const int* __restrict__ indices, int n_rays, int S, float* __restrict__ grad_table)
{
    // samples == null: positions are rebuilt from the packed hit records with the sampler's arithmetic
    const float inv_s = 1.0f / (float)S;
    const bool pow2 = (S & (S - 1)) == 0;
    const int log2s = 31 - __clz(S);
    const unsigned FULL = 0xffffffffu;
    __shared__ uint32_t stage[HB_WARPS * 32][NET_IN / 2 + 1];   // +1: conflict-free per-thread rows
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int n_groups = (n_rays + 31) >> 5;
    for (int g = blockIdx.x; g < n_groups; g += gridDim.x) {
        const int ray = g * 32 + lane;
        int ns = 0;
        long long base = 0;
        if (ray < n_rays) {
            ns = num_hits[ray] * S;
            base = (long long)indices[ray] * S;
        }
        int maxn = ns;
#pragma unroll
        for (int o = 16; o; o >>= 1) maxn = max(maxn, __shfl_xor_sync(FULL, maxn, o));
        for (int q = warp; q < maxn; q += HB_WARPS) {     // warp-uniform trip count
            const bool valid = q < ns;
            const long long s = base + q;
            float x01 = 0.5f, y01 = 0.5f, z01 = 0.5f;
            // the row's 16 feature-gradient pairs: four 16-byte loads staged in shared memory (the
            // level loop is rolled; a 4-byte load per level would cost 32 L1TEX wavefronts each)
            uint32_t* my = stage[threadIdx.x];
            if (valid) {
                if (samples) {
                    const float* p = samples + s * stride;
                    x01 = unit_coord(ld_stream_f1(p));
                    y01 = unit_coord(ld_stream_f1(p + 1));
                    z01 = unit_coord(ld_stream_f1(p + 2));
                } else {                                     // sampler.cu:45-59, REGULAR
                    const int j = pow2 ? (q >> log2s) : (q / S);
                    const int i = q - j * S;
                    uint32_t rec[8];
                    ldg_v8(seg8 + 8 * (base / S + j), rec);  // base = indices[ray] * S
                    const float sx = __uint_as_float(rec[0]), sy = __uint_as_float(rec[1]), sz = __uint_as_float(rec[2]);
                    float tt = 0.0f;
                    if (pow2) tt = rn_mul((float)i, inv_s);
                    else
                        for (int k = 0; k < i; k++) tt = rn_add(tt, inv_s);
                    x01 = unit_coord(rn_fma(tt, rn_sub(__uint_as_float(rec[3]), sx), sx));
                    y01 = unit_coord(rn_fma(tt, rn_sub(__uint_as_float(rec[4]), sy), sy));
                    z01 = unit_coord(rn_fma(tt, rn_sub(__uint_as_float(rec[5]), sz), sz));
                }
                // the 64-byte row in two 256-bit loads (the lanes' rows lie a ray apart: one tag lookup per lane and
                // instruction, whatever the width)
                const uint16_t* ge = d_encoded + s * NET_IN;
#pragma unroll
                for (int c = 0; c < 2; c++) {
                    uint32_t e[8];
                    ldg_v8(ge + 16 * c, e);
#pragma unroll
                    for (int i = 0; i < 8; i++) my[8 * c + i] = e[i];
                }
            }
#pragma unroll 1
            for (int l = 0; l < lv.n_levels; l++) {
                float2 gl = make_float2(0.f, 0.f);
                if (valid) gl = h2_bits_to_f2(my[l]);
                const bool live = gl.x != 0.0f || gl.y != 0.0f;
                if (!__any_sync(FULL, live)) continue;
                LevelPoint pt;
                level_point(lv, l, x01, y01, z01, pt);
                // cell key; dead lanes get a key of their own so that they end a run
                const uint32_t k0 = live ? (pt.gx | (pt.gy << 16)) : 0xffffffffu;
                const uint32_t k1 = live ? pt.gz : 0xffffffffu;
                const uint32_t p0 = __shfl_up_sync(FULL, k0, 1), p1 = __shfl_up_sync(FULL, k1, 1);
                const bool head = lane == 0 || p0 != k0 || p1 != k1;
                const unsigned heads = __ballot_sync(FULL, head);
                float v[16];
#pragma unroll
                for (int c = 0; c < 8; c++) {
                    v[2 * c] = pt.w[c] * gl.x;
                    v[2 * c + 1] = pt.w[c] * gl.y;
                }
                if (heads == 1u) {
                    // the whole warp sits in ONE cell (coarse levels): reduce-scatter instead of 5 x 16
                    // butterflies.  Two halving rounds leave each lane one x-pair quad (4 values), three
                    // plain rounds finish it: 24 shuffles instead of 80, and ONE 16-byte reduction per quad.
                    const bool b4 = lane & 16, b3 = lane & 8;
                    float h8[8];
#pragma unroll
                    for (int i = 0; i < 8; i++) {
                        const float mine = b4 ? v[8 + i] : v[i], send = b4 ? v[i] : v[8 + i];
                        h8[i] = mine + __shfl_xor_sync(FULL, send, 16);
                    }
                    float q4[4];
#pragma unroll
                    for (int i = 0; i < 4; i++) {
                        const float mine = b3 ? h8[4 + i] : h8[i], send = b3 ? h8[i] : h8[4 + i];
                        q4[i] = mine + __shfl_xor_sync(FULL, send, 8);
                    }
#pragma unroll
                    for (int d = 4; d; d >>= 1)
#pragma unroll
                        for (int i = 0; i < 4; i++) q4[i] += __shfl_xor_sync(FULL, q4[i], d);
                    if ((lane & 7) == 0) {       // lanes 0, 8, 16, 24 hold quads 0, 1, 2, 3 (qd = 2*b4 + b3)
                        const int qd = lane >> 3;
                        const uint32_t i0 = qd == 0 ? pt.idx[0] : qd == 1 ? pt.idx[2] : qd == 2 ? pt.idx[4] : pt.idx[6];
                        const uint32_t i1 = qd == 0 ? pt.idx[1] : qd == 1 ? pt.idx[3] : qd == 2 ? pt.idx[5] : pt.idx[7];
                        float2* gt = reinterpret_cast<float2*>(grad_table) + lv.offset[l];
                        const float2 a = make_float2(q4[0], q4[1]), b = make_float2(q4[2], q4[3]);
                        if ((i0 ^ i1) == 1u) {
                            const bool odd = i0 & 1u;
                            atomicAdd(reinterpret_cast<float4*>(gt + (i0 & ~1u)),
                                      odd ? make_float4(b.x, b.y, a.x, a.y) : make_float4(a.x, a.y, b.x, b.y));
                        } else {
                            atomicAdd(gt + i0, a);
                            atomicAdd(gt + i1, b);
                        }
                    }
                    continue;
                }
                // lanes i and i+d are in the same run iff no head in (i, i+d]
                const unsigned after = lane < 31 ? (heads >> (lane + 1)) : 1u;
#pragma unroll
                for (int d = 1; d < 32; d <<= 1) {
                    const bool same = (lane + d < 32) && (after & ((1u << d) - 1u)) == 0u;
                    if (!__any_sync(FULL, same)) break;     // longest run reached
#pragma unroll
                    for (int i = 0; i < 16; i++) {
                        const float t = __shfl_down_sync(FULL, v[i], d);
                        if (same) v[i] += t;
                    }
                }
                if (head && live) {
                    float2* gt = reinterpret_cast<float2*>(grad_table) + lv.offset[l];
#pragma unroll
                    for (int qd = 0; qd < 4; qd++) {
                        const uint32_t i0 = pt.idx[2 * qd], i1 = pt.idx[2 * qd + 1];
                        const float2 a = make_float2(v[4 * qd], v[4 * qd + 1]), b = make_float2(v[4 * qd + 2], v[4 * qd + 3]);
                        if ((i0 ^ i1) == 1u) {
                            const bool odd = i0 & 1u;
                            atomicAdd(reinterpret_cast<float4*>(gt + (i0 & ~1u)),
                                      odd ? make_float4(b.x, b.y, a.x, a.y) : make_float4(a.x, a.y, b.x, b.y));
                        } else {
                            atomicAdd(gt + i0, a);
                            atomicAdd(gt + i1, b);
                        }
                    }
                }
            }
        }
    }
}

// ---------------------------------------------------------------------------------------
// L2 loss (tiny-cuda-nn loss->evaluate as called at main.cu:755-759, restated): per element
//   diff = pred - target;  values = diff^2 / n;  gradients = loss_scale * 2 * diff / n  (fp16)
// with n = batch_size * 3.  The prediction is taken in fp32 (the pixels buffer of
// launch_volrender_cuda) — the intended wiring of Q4 without the half round trip.
// ---------------------------------------------------------------------------------------
__global__ void l2_loss_kernel(const float* __restrict__ pred, const float* __restrict__ target, long long n,
                               float loss_scale, float* __restrict__ values, __half* __restrict__ gradients,
                               float* __restrict__ loss_sum)
{
    const float inv_n = 1.0f / (float)n;
    float local = 0.0f;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const float diff = pred[i] - target[i];
        const float v = diff * diff * inv_n;
        if (values) values[i] = v;
        gradients[i] = __float2half_rn(loss_scale * 2.0f * diff * inv_n);
        local += v;
    }
    if (loss_sum) {   // tcnn::reduce_sum(values) (main.cu:760)
#pragma unroll
        for (int o = 16; o; o >>= 1) local += __shfl_xor_sync(0xffffffffu, local, o);
        if ((threadIdx.x & 31) == 0) atomicAdd(loss_sum, local);
    }
}

// ---------------------------------------------------------------------------------------
// Adam (tiny-cuda-nn optimizer->step as called at main.cu:787 with the config of main.cu:40-46:
// lr 1e-3, beta1 0.9, beta2 0.999, eps 1e-8, no L2 reg), restated; parameter layout as the
// driver's (main.cu:328-342): fp32 master, fp16 inference copy, gradient.  Gradients here are
// fp32 (already summed over ranks) and are ZEROED on the way out, so no separate memset.
//   g = grad * grad_scale (grad_scale = 1 / (loss_scale * world)) ; m,v moments ;
//   lr_t = lr * sqrt(1 - beta2^t) / (1 - beta1^t) ; master -= lr_t * m / (sqrt(v) + eps) ; half = fp16(master)
// ---------------------------------------------------------------------------------------
// One element, spelled with the _rn intrinsics so that nvcc's FMA contraction cannot differ between the
// two kernels below (the sequence is the one the plain expressions compiled to: fma(g, 1-b1, b1*m1),
// fma(g, g*(1-b2), b2*m2)).
__device__ __forceinline__ float adam_one(float g_raw, float& m1, float& m2, float w, float grad_scale, float lr_t,
                                          float beta1, float beta2, float eps)
{
    const float g = __fmul_rn(g_raw, grad_scale);
    const float a = __fmaf_rn(g, __fsub_rn(1.0f, beta1), __fmul_rn(beta1, m1));
    const float b = __fmaf_rn(g, __fmul_rn(g, __fsub_rn(1.0f, beta2)), __fmul_rn(beta2, m2));
    m1 = a;
    m2 = b;
    return __fsub_rn(w, __fdiv_rn(__fmul_rn(lr_t, a), __fadd_rn(__fsqrt_rn(b), eps)));
}

// n_skip: the first n_skip elements (of this launch) are ENCODING parameters, which tiny-cuda-nn's Adam leaves
// untouched — weight and both moments — in a step where their gradient is exactly zero (a hash-table entry no
// sample of the batch touched); matrix weights always take the update.
__global__ void adam_step_kernel(long long n, long long n_skip, float* __restrict__ master, __half* __restrict__ params,
                                 float* __restrict__ grads, float* __restrict__ m1, float* __restrict__ m2,
                                 float grad_scale, float lr_t, float beta1, float beta2, float eps)
{
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        float a = m1[i], b = m2[i];
        const float g = grads[i];
        if (i < n_skip && g == 0.0f) continue;
        const float w = adam_one(g, a, b, master[i], grad_scale, lr_t, beta1, beta2, eps);
        grads[i] = 0.0f;
        m1[i] = a;
        m2[i] = b;
        master[i] = w;
        params[i] = __float2half_rn(w);
    }
}

// Four parameters per thread, every load issued before the first use, streaming hints (each
// array is touched exactly once per step): the scalar loop above ran at 0.7 TB/s (0.62 ms for 12.2 M
// parameters; every warp stalled on its loads, DRAM at 7 %), this one takes 0.19 ms.  Same adam_one
// per element, so the same bits.
__device__ __forceinline__ void adam_four(long long i, long long n_skip, const float4& g, float4& a, float4& b, float4& w,
                                          float4* __restrict__ master, uint2* __restrict__ params,
                                          float4* __restrict__ grads, float4* __restrict__ m1, float4* __restrict__ m2,
                                          float grad_scale, float lr_t, float beta1, float beta2, float eps)
{
    const long long e = 4 * i;       // first element of this group; elements below n_skip skip on a zero gradient
    if (e + 3 < n_skip && g.x == 0.0f && g.y == 0.0f && g.z == 0.0f && g.w == 0.0f) return;   // nothing changes
    __stcs(grads + i, make_float4(0.0f, 0.0f, 0.0f, 0.0f));
    if (!(e < n_skip && g.x == 0.0f)) w.x = adam_one(g.x, a.x, b.x, w.x, grad_scale, lr_t, beta1, beta2, eps);
    if (!(e + 1 < n_skip && g.y == 0.0f)) w.y = adam_one(g.y, a.y, b.y, w.y, grad_scale, lr_t, beta1, beta2, eps);
    if (!(e + 2 < n_skip && g.z == 0.0f)) w.z = adam_one(g.z, a.z, b.z, w.z, grad_scale, lr_t, beta1, beta2, eps);
    if (!(e + 3 < n_skip && g.w == 0.0f)) w.w = adam_one(g.w, a.w, b.w, w.w, grad_scale, lr_t, beta1, beta2, eps);
    __stcs(m1 + i, a);
    __stcs(m2 + i, b);
    __stcs(master + i, w);
    const __half2 lo = __floats2half2_rn(w.x, w.y), hi = __floats2half2_rn(w.z, w.w);
    params[i] = make_uint2(*reinterpret_cast<const unsigned int*>(&lo), *reinterpret_cast<const unsigned int*>(&hi));
}

__global__ void __launch_bounds__(256)
adam_step_vec4_kernel(long long n4, long long n_skip, float4* __restrict__ master, uint2* __restrict__ params, float4* __restrict__ grads,
                      float4* __restrict__ m1, float4* __restrict__ m2, float grad_scale, float lr_t, float beta1,
                      float beta2, float eps)
{
    const long long stride = (long long)gridDim.x * blockDim.x;
    // two float4 groups per iteration: eight 16-byte loads in flight per thread before the first use
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += 2 * stride) {
        const long long j = i + stride;
        const bool two = j < n4;
        const float4 z = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
        const float4 g0 = __ldcs(grads + i);
        float4 a0 = __ldcs(m1 + i);
        float4 b0 = __ldcs(m2 + i);
        float4 w0 = __ldcs(master + i);
        const float4 g1 = two ? __ldcs(grads + j) : z;
        float4 a1 = two ? __ldcs(m1 + j) : z;
        float4 b1 = two ? __ldcs(m2 + j) : z;
        float4 w1 = two ? __ldcs(master + j) : z;
        adam_four(i, n_skip, g0, a0, b0, w0, master, params, grads, m1, m2, grad_scale, lr_t, beta1, beta2, eps);
        if (two) adam_four(j, n_skip, g1, a1, b1, w1, master, params, grads, m1, m2, grad_scale, lr_t, beta1, beta2, eps);
    }
}

// ------------------------------------------------------------------------------------ host ABI
// The two stages of the network backward as separate entry points: a data-parallel caller issues the all-reduce
// of the (28 KB) MLP weight gradients as soon as the first has finished, under the scatter of the second.
extern "C" int rtxnerf_mlp_backward(const rtxnerf_mlp* mlp, const uint16_t* d_encoded, const uint16_t* d_dl_dradiance,
                                    const long long* d_count, long long count_scale, long long n_max,
                                    uint16_t* d_dencoded, float* d_grad_weights, void* stream)
{
    if (!mlp || !mlp->weights || !d_encoded || !d_dl_dradiance || !d_dencoded || !d_grad_weights) return RTXNERF_EINVAL;
    if (n_max < 0 || count_scale < 1) return RTXNERF_EINVAL;
    if (mlp->in_dim != NET_IN || mlp->width != NET_W || mlp->n_hidden != 2) return RTXNERF_ECONFIG;
    if (mlp->out_activation != 0 && mlp->out_activation != 1) return RTXNERF_ECONFIG;
    if (n_max == 0) return RTXNERF_OK;
    const long long tiles = (n_max + NET_TILE - 1) / NET_TILE;
    // RTXNERF_MLPB_VERSION=1 keeps the one-tile-per-CTA kernel (development A/B)
    static const int version = rtx_env_int("RTXNERF_MLPB_VERSION", 2);
    if (version == 2) {
        static RtxPerDeviceOnce attr2;
        RTX_CUDA_TRY((cudaError_t)rtx_set_max_smem(attr2, mlp_backward2_kernel, N2Smem::BYTES));
        const long long ctas = (tiles + N2_GROUPS - 1) / N2_GROUPS;
        const long long res2 = (long long)RTX_NUM_SMS * 2;          // 2 CTAs per SM: 2 x 256 TMEM columns, 2 x 102 KB
        mlp_backward2_kernel<<<(unsigned)(ctas < res2 ? ctas : res2), N2_THREADS, N2Smem::BYTES, (cudaStream_t)stream>>>(
            mlp->weights, d_encoded, d_dl_dradiance, mlp->out_activation, d_count, count_scale, n_max, d_dencoded,
            d_grad_weights);
        return rtx_launch_status();
    }
    static RtxPerDeviceOnce attr_set;
    RTX_CUDA_TRY((cudaError_t)rtx_set_max_smem(attr_set, mlp_backward_kernel, NbSmem::BYTES));
    const long long resident = (long long)RTX_NUM_SMS * 2;   // 2 CTAs per SM: 2 x 256 TMEM columns, 2 x 75 KB
    mlp_backward_kernel<<<(unsigned)(tiles < resident ? tiles : resident), NET_TILE, NbSmem::BYTES, (cudaStream_t)stream>>>(
        mlp->weights, d_encoded, d_dl_dradiance, mlp->out_activation, d_count, count_scale, n_max, d_dencoded,
        d_grad_weights);
    return rtx_launch_status();
}

extern "C" int rtxnerf_hashgrid_backward_rays(const rtxnerf_hashgrid* enc, const float* d_samples, int sample_stride,
                                              const uint16_t* d_dencoded, const int* d_num_hits, const int* d_indices,
                                              int n_rays, int samples_per_segment, float* d_grad_table, void* stream)
{
    if (!enc || !d_samples || !d_dencoded || !d_num_hits || !d_indices || !d_grad_table) return RTXNERF_EINVAL;
    if (n_rays < 0 || samples_per_segment < 1 || sample_stride < 3) return RTXNERF_EINVAL;
    if ((uintptr_t)d_dencoded & 31u) return RTXNERF_EINVAL;         // 64-byte rows read with 256-bit loads
    Levels lv;
    int rc = levels_from(enc, &lv);
    if (rc) return rc;
    if (enc->n_levels * 2 > NET_IN) return RTXNERF_ECONFIG;
    if (n_rays == 0) return RTXNERF_OK;
    const long long groups = ((long long)n_rays + 31) / 32;
    // one CTA per 32-ray group up to 64 per SM: the hardware scheduler balances the groups (their sample counts
    // differ); a static grid-stride split over 8 CTAs per SM measured 2 % slower (10.61 -> 10.38 ms at 2^18 rays)
#ifndef HB_BLOCKS_PER_SM
#define HB_BLOCKS_PER_SM 64
#endif
    const long long max_blocks = (long long)RTX_NUM_SMS * HB_BLOCKS_PER_SM;
    hashgrid_backward_kernel<<<(unsigned)(groups < max_blocks ? groups : max_blocks), HB_WARPS * 32, 0,
                               (cudaStream_t)stream>>>(lv, d_samples, sample_stride, nullptr, d_dencoded, d_num_hits,
                                                       d_indices, n_rays, samples_per_segment, d_grad_table);
    return rtx_launch_status();
}

// The scatter with the sample positions rebuilt from the packed hit records (REGULAR sampling) instead of read
// from a sample buffer: the backward half of rtxnerf_network_forward_seg.
extern "C" int rtxnerf_hashgrid_backward_seg(const rtxnerf_hashgrid* enc, const float* d_segments8,
                                             const uint16_t* d_dencoded, const int* d_num_hits, const int* d_indices,
                                             int n_rays, int samples_per_segment, float* d_grad_table, void* stream)
{
    if (!enc || !d_segments8 || !d_dencoded || !d_num_hits || !d_indices || !d_grad_table) return RTXNERF_EINVAL;
    if (n_rays < 0 || samples_per_segment < 1) return RTXNERF_EINVAL;
    if (((uintptr_t)d_segments8 & 31u) || ((uintptr_t)d_dencoded & 31u)) return RTXNERF_EINVAL;
    Levels lv;
    int rc = levels_from(enc, &lv);
    if (rc) return rc;
    if (enc->n_levels * 2 > NET_IN) return RTXNERF_ECONFIG;
    if (n_rays == 0) return RTXNERF_OK;
    const long long groups = ((long long)n_rays + 31) / 32;
    const long long max_blocks = (long long)RTX_NUM_SMS * HB_BLOCKS_PER_SM;
    hashgrid_backward_kernel<<<(unsigned)(groups < max_blocks ? groups : max_blocks), HB_WARPS * 32, 0,
                               (cudaStream_t)stream>>>(lv, nullptr, 0, d_segments8, d_dencoded, d_num_hits, d_indices,
                                                       n_rays, samples_per_segment, d_grad_table);
    return rtx_launch_status();
}

extern "C" int rtxnerf_network_backward_rays(const rtxnerf_hashgrid* enc, const rtxnerf_mlp* mlp,
                                             const float* d_samples, int sample_stride, const uint16_t* d_encoded,
                                             const uint16_t* d_dl_dradiance, const int* d_num_hits,
                                             const int* d_indices, int n_rays, int samples_per_segment,
                                             const long long* d_count, long long n_max, uint16_t* d_dencoded,
                                             float* d_grad_table, float* d_grad_weights, void* stream)
{
    if (!enc || !mlp || !d_samples || !d_num_hits || !d_indices || !d_grad_table) return RTXNERF_EINVAL;
    if (n_rays < 0 || samples_per_segment < 1 || sample_stride < 3 || n_max < 0) return RTXNERF_EINVAL;
    if (n_rays == 0 || n_max == 0) {      // argument errors are still reported for an empty batch
        Levels lv;
        if (!mlp->weights || !d_encoded || !d_dl_dradiance || !d_dencoded || !d_grad_weights) return RTXNERF_EINVAL;
        if (mlp->in_dim != NET_IN || mlp->width != NET_W || mlp->n_hidden != 2) return RTXNERF_ECONFIG;
        return levels_from(enc, &lv);
    }
    int rc = rtxnerf_mlp_backward(mlp, d_encoded, d_dl_dradiance, d_count, samples_per_segment, n_max, d_dencoded,
                                  d_grad_weights, stream);
    if (rc) return rc;
    return rtxnerf_hashgrid_backward_rays(enc, d_samples, sample_stride, d_dencoded, d_num_hits, d_indices, n_rays,
                                          samples_per_segment, d_grad_table, stream);
}

extern "C" int rtxnerf_l2_loss(const float* d_pred, const float* d_target, long long n, float loss_scale,
                               float* d_values, uint16_t* d_gradients, float* d_loss_sum, void* stream)
{
    if (!d_pred || !d_target || !d_gradients || n < 0) return RTXNERF_EINVAL;
    if (n == 0) return RTXNERF_OK;
    long long blocks = (n + 255) / 256;
    if (blocks > RTX_NUM_SMS * 8) blocks = RTX_NUM_SMS * 8;
    l2_loss_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(d_pred, d_target, n, loss_scale, d_values,
                                                                      reinterpret_cast<__half*>(d_gradients), d_loss_sum);
    return rtx_launch_status();
}

extern "C" int rtxnerf_adam_step(long long n, float* d_master, uint16_t* d_params, float* d_grads, float* d_m1,
                                 float* d_m2, float grad_scale, float lr, float beta1, float beta2, float eps,
                                 int step, void* stream)
{
    return rtxnerf_adam_step_sparse(n, 0, d_master, d_params, d_grads, d_m1, d_m2, grad_scale, lr, beta1, beta2, eps,
                                    step, stream);
}

extern "C" int rtxnerf_adam_step_sparse(long long n, long long n_encoding, float* d_master, uint16_t* d_params,
                                        float* d_grads, float* d_m1, float* d_m2, float grad_scale, float lr,
                                        float beta1, float beta2, float eps, int step, void* stream)
{
    if (!d_master || !d_params || !d_grads || !d_m1 || !d_m2 || n < 0 || step < 1) return RTXNERF_EINVAL;
    if (n_encoding < 0 || n_encoding > n) return RTXNERF_EINVAL;
    if (n == 0) return RTXNERF_OK;
    const float lr_t = lr * sqrtf(1.0f - powf(beta2, (float)step)) / (1.0f - powf(beta1, (float)step));
    static const int adam_scalar = rtx_env_int("RTXNERF_ADAM_SCALAR", 0);   // development knob
    // body: 4 parameters per thread when every array is 16-byte aligned (8 for the fp16 copy); the tail
    // (n % 4) and unaligned callers go through the scalar kernel
    const bool aligned = (((uintptr_t)d_master | (uintptr_t)d_grads | (uintptr_t)d_m1 | (uintptr_t)d_m2) & 15) == 0 &&
                         ((uintptr_t)d_params & 7) == 0 && !adam_scalar;
    const long long n4 = aligned ? n / 4 : 0;
    int launched = 0;
    if (n4 > 0) {
        long long blocks = (n4 + 255) / 256;
        if (blocks > RTX_NUM_SMS * 8) blocks = RTX_NUM_SMS * 8;
        adam_step_vec4_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(
            n4, n_encoding, reinterpret_cast<float4*>(d_master), reinterpret_cast<uint2*>(d_params), reinterpret_cast<float4*>(d_grads),
            reinterpret_cast<float4*>(d_m1), reinterpret_cast<float4*>(d_m2), grad_scale, lr_t, beta1, beta2, eps);
        launched++;
    }
    const long long done = n4 * 4, rest = n - done;
    if (rest > 0) {
        long long blocks = (rest + 255) / 256;
        if (blocks > RTX_NUM_SMS * 16) blocks = RTX_NUM_SMS * 16;
        adam_step_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(
            rest, n_encoding > done ? n_encoding - done : 0, d_master + done, reinterpret_cast<__half*>(d_params) + done, d_grads + done, d_m1 + done, d_m2 + done,
            grad_scale, lr_t, beta1, beta2, eps);
        launched++;
    }
    return rtx_launch_status(launched);
}
