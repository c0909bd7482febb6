// network.cu — multiresolution hash encoding + fully fused MLP forward for sm_100a.
//
// Replaces tiny-cuda-nn's encoding + FullyFusedMLP forward as the reference calls it
// (network->forward, main.cu:721; output slice + convertHalfToFloat, main.cu:723-728).
// tcnn runs 16x16x16 wmma fragments (HMMA, sm_70 generation); here the three dense layers
// 32->64->64->16 run on the 5th-gen tensor cores:
//   * one CTA = 128 threads owns a tile of 128 samples = the M dimension of tcgen05.mma
//   * weights (14 KB fp16) live in shared memory for the CTA's lifetime, in the K-major
//     no-swizzle core-matrix layout the UMMA descriptors address
//   * each thread encodes ITS sample (16 levels x 8 corner gathers, table L2-resident) and
//     writes its row of the A operand straight into the UMMA layout
//   * accumulators live in TMEM (128 lanes x 64 fp32 columns); epilogue = tcgen05.ld ->
//     ReLU -> fp16 -> next layer's A operand in shared memory (activations never touch HBM)
//   * last epilogue: sigmoid, fp16 round, and the fp32 (r,g,b,sigma) gather fused in, so the
//     reference's separate half->float kernel (K7) disappears
// Persistent grid; several CTAs per SM interleave gather latency with MMA/epilogue work.
#include "common.cuh"
#include "hashgrid.cuh"
#include "mlp_tile.cuh"

// encoding only (parity check of the features; also the producer for the SIMT check kernel)
__global__ void __launch_bounds__(256)
hashgrid_encode_kernel(Levels lv, const __half2* __restrict__ table, const float* __restrict__ samples, int stride,
                       const long long* __restrict__ d_count, long long count_scale, long long n_max,
                       __half2* __restrict__ encoded)
{
    const long long n = row_count(d_count, count_scale, n_max);
    for (long long s = (long long)blockIdx.x * blockDim.x + threadIdx.x; s < n; s += (long long)gridDim.x * blockDim.x) {
        const float* p = samples + s * stride;
        const float x01 = unit_coord(p[0]), y01 = unit_coord(p[1]), z01 = unit_coord(p[2]);
        for (int l = 0; l < lv.n_levels; l++)
            encoded[s * lv.n_levels + l] =
                encode_level(table, lv, l, x01, y01, z01);
    }
}

struct NetSmem {
    // offsets into dynamic shared memory (all multiples of 128).  One activation tile: the A
    // operand of layer k+1 overwrites that of layer k, which is safe because every thread has
    // waited on the MMA-complete barrier of layer k before it writes.  30 KB per CTA -> 7 CTAs
    // per SM, 7 x 64 = 448 of the 512 TMEM columns.
    static constexpr int W0 = 0;                       // [64][32]  4 KB
    static constexpr int W1 = W0 + NET_W * NET_IN * 2; // [64][64]  8 KB
    static constexpr int W2 = W1 + NET_W * NET_W * 2;  // [16][64]  2 KB
    static constexpr int ACT = W2 + NET_OUT * NET_W * 2;   // [128][32] then [128][64] twice, 16 KB
    static constexpr int BAR = ACT + NET_TILE * NET_W * 2; // mbarrier (8 B) + tmem address (4 B)
    static constexpr int BYTES = BAR + 16;
};
#define NET_CTAS_PER_SM 7

template <bool ENCODE>
__global__ void __launch_bounds__(NET_TILE)
network_forward_tc_kernel(Levels lv, const __half2* __restrict__ table, const uint16_t* __restrict__ weights,
                          const void* __restrict__ input, int input_stride, int out_activation,
                          const long long* __restrict__ d_count, long long count_scale, long long n_max,
                          uint16_t* __restrict__ out16, float* __restrict__ radiance4)
{
    extern __shared__ __align__(128) uint8_t smem[];
    uint64_t* bar = reinterpret_cast<uint64_t*>(smem + NetSmem::BAR);
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + NetSmem::BAR + 8);
    const int tid = threadIdx.x, warp = tid >> 5;

    load_weight_tile(smem + NetSmem::W0, weights, NET_W, NET_IN, tid, NET_TILE);
    load_weight_tile(smem + NetSmem::W1, weights + NET_W * NET_IN, NET_W, NET_W, tid, NET_TILE);
    load_weight_tile(smem + NetSmem::W2, weights + NET_W * NET_IN + NET_W * NET_W, NET_OUT, NET_W, tid, NET_TILE);
    if (tid == 0) {
        tc::mbar_init(bar, 1);
        tc::fence_mbar_init();
    }
    if (warp == 0) tc::tmem_alloc<NET_TMEM_COLS>(tmem_slot);
    tc::fence_proxy_async_smem();
    tc::fence_before_thread_sync();
    __syncthreads();
    tc::fence_after_thread_sync();
    const uint32_t tmem_base = *tmem_slot;
    const uint32_t t_row = tmem_base + ((uint32_t)(warp * 32) << 16);   // this warp's 32 TMEM lanes
    const uint32_t s_w0 = tc::smem_u32(smem + NetSmem::W0), s_w1 = tc::smem_u32(smem + NetSmem::W1),
                   s_w2 = tc::smem_u32(smem + NetSmem::W2), s_act = tc::smem_u32(smem + NetSmem::ACT);
    uint8_t* act = smem + NetSmem::ACT;

    const long long n = row_count(d_count, count_scale, n_max);
    const long long n_tiles = (n + NET_TILE - 1) / NET_TILE;
    uint32_t phase = 0;
    uint4 pre[NET_IN / 8];   // software pipeline of the pre-encoded input rows (ENCODE == false)
    if (!ENCODE) {
        const long long s0 = (long long)blockIdx.x * NET_TILE + tid;
        const uint4* src = reinterpret_cast<const uint4*>(reinterpret_cast<const uint16_t*>(input) + s0 * NET_IN);
#pragma unroll
        for (int c = 0; c < NET_IN / 8; c++) pre[c] = s0 < n ? __ldg(src + c) : make_uint4(0, 0, 0, 0);
    }
    for (long long tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        const long long s = tile * NET_TILE + tid;
        const bool valid = s < n;
        // ---- A0: this thread's row of the layer-0 operand.  The previous tile's last MMA has
        // completed (every thread waited on its barrier), so the tile may be overwritten.
        if (ENCODE) {
            float x01 = 0.5f, y01 = 0.5f, z01 = 0.5f;
            if (valid) {
                const float* p = reinterpret_cast<const float*>(input) + s * input_stride;
                x01 = unit_coord(ld_stream_f1(p));
                y01 = unit_coord(ld_stream_f1(p + 1));
                z01 = unit_coord(ld_stream_f1(p + 2));
            }
            // rolled over levels: per-level parameters come from the constant bank, the body
            // stays small enough for the instruction cache, 8 gathers are in flight per thread
            // and the other CTAs of the SM hide their latency
#pragma unroll 1
            for (int l = 0; l < NET_IN / 2; l++) {
                __half2 f = __floats2half2_rn(0.0f, 0.0f);
                if (l < lv.n_levels) f = encode_level(table, lv, l, x01, y01, z01);
                *reinterpret_cast<__half2*>(act + tile_chunk_off(tid, l >> 2, NET_IN) + (l & 3) * 4) = f;
            }
        } else {
            // this tile's row was fetched while the previous tile ran its layers; fetch the next one now
#pragma unroll
            for (int c = 0; c < NET_IN / 8; c++)
                *reinterpret_cast<uint4*>(act + tile_chunk_off(tid, c, NET_IN)) = pre[c];
            const long long sn = (tile + gridDim.x) * NET_TILE + tid;
            const uint4* src = reinterpret_cast<const uint4*>(reinterpret_cast<const uint16_t*>(input) + sn * NET_IN);
#pragma unroll
            for (int c = 0; c < NET_IN / 8; c++) pre[c] = sn < n ? __ldg(src + c) : make_uint4(0, 0, 0, 0);
        }
        tc::fence_proxy_async_smem();
        tc::fence_before_thread_sync();   // also orders the previous tile's tcgen05.ld before the next MMA
        __syncthreads();
        // ---- layer 0: D = A0 * W0^T
        if (tid == 0) {
            tc::fence_after_thread_sync();
            issue_layer<NET_W, NET_IN>(tmem_base, s_act, s_w0, bar);
        }
        tc::mbar_wait(bar, phase);
        phase ^= 1;
        tc::fence_after_thread_sync();
        relu_row_to_smem(t_row, act, tid);
        tc::fence_proxy_async_smem();
        tc::fence_before_thread_sync();
        __syncthreads();
        // ---- layer 1: D = A1 * W1^T
        if (tid == 0) {
            tc::fence_after_thread_sync();
            issue_layer<NET_W, NET_W>(tmem_base, s_act, s_w1, bar);
        }
        tc::mbar_wait(bar, phase);
        phase ^= 1;
        tc::fence_after_thread_sync();
        relu_row_to_smem(t_row, act, tid);
        tc::fence_proxy_async_smem();
        tc::fence_before_thread_sync();
        __syncthreads();
        // ---- output layer: D[:, 0:16] = A2 * W2^T
        if (tid == 0) {
            tc::fence_after_thread_sync();
            issue_layer<NET_OUT, NET_W>(tmem_base, s_act, s_w2, bar);
        }
        tc::mbar_wait(bar, phase);
        phase ^= 1;
        tc::fence_after_thread_sync();
        uint32_t r[16];
        output_row_load(t_row, r);
        if (valid) output_row_store(r, out_activation, s, out16, radiance4);
    }
    tc::fence_before_thread_sync();
    __syncthreads();
    tc::fence_after_thread_sync();
    if (warp == 0) tc::tmem_dealloc<NET_TMEM_COLS>(tmem_base);
}

// ---------------------------------------------------------------------------------------
// CUDA-core check kernel (same numerics contract: fp16 weights/activations, fp32 accumulate).
// Not the product path: selected only with impl = RTXNERF_IMPL_SIMT for A/B checks on the GPU.
// ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128)
network_forward_simt_kernel(Levels lv, const __half2* __restrict__ table, const uint16_t* __restrict__ weights,
                            const void* __restrict__ input, int input_stride, int encode, int out_activation,
                            const long long* __restrict__ d_count, long long count_scale, long long n_max,
                            uint16_t* __restrict__ out16, float* __restrict__ radiance4)
{
    __shared__ __half w[NET_W * NET_IN + NET_W * NET_W + NET_OUT * NET_W];
    for (int i = threadIdx.x; i < NET_W * NET_IN + NET_W * NET_W + NET_OUT * NET_W; i += blockDim.x)
        w[i] = reinterpret_cast<const __half*>(weights)[i];
    __syncthreads();
    const long long n = row_count(d_count, count_scale, n_max);
    for (long long s = (long long)blockIdx.x * blockDim.x + threadIdx.x; s < n; s += (long long)gridDim.x * blockDim.x) {
        float a[NET_W], b[NET_W];
        if (encode) {
            const float* p = reinterpret_cast<const float*>(input) + s * input_stride;
            const float x01 = unit_coord(p[0]), y01 = unit_coord(p[1]), z01 = unit_coord(p[2]);
            for (int l = 0; l < NET_IN / 2; l++) {
                float2 f = make_float2(0.f, 0.f);
                if (l < lv.n_levels)
                    f = __half22float2(encode_level(table, lv, l, x01, y01, z01));
                a[2 * l] = f.x;
                a[2 * l + 1] = f.y;
            }
        } else {
            const __half* src = reinterpret_cast<const __half*>(input) + s * NET_IN;
            for (int k = 0; k < NET_IN; k++) a[k] = __half2float(src[k]);
        }
        const __half* wl = w;
        for (int r = 0; r < NET_W; r++) {
            float acc = 0.f;
            for (int k = 0; k < NET_IN; k++) acc = fmaf(__half2float(wl[r * NET_IN + k]), a[k], acc);
            b[r] = __half2float(__float2half_rn(fmaxf(acc, 0.f)));
        }
        wl += NET_W * NET_IN;
        for (int r = 0; r < NET_W; r++) {
            float acc = 0.f;
            for (int k = 0; k < NET_W; k++) acc = fmaf(__half2float(wl[r * NET_W + k]), b[k], acc);
            a[r] = __half2float(__float2half_rn(fmaxf(acc, 0.f)));
        }
        wl += NET_W * NET_W;
        for (int r = 0; r < NET_OUT; r++) {
            float acc = 0.f;
            for (int k = 0; k < NET_W; k++) acc = fmaf(__half2float(wl[r * NET_W + k]), a[k], acc);
            if (out_activation == 1) acc = 1.0f / (1.0f + expf(-acc));
            const __half h = __float2half_rn(acc);
            if (out16) reinterpret_cast<__half*>(out16)[s * NET_OUT + r] = h;
            if (radiance4 && r < 4) radiance4[s * 4 + r] = __half2float(h);
        }
    }
}

// ------------------------------------------------------------------------------------ host ABI
extern "C" int rtxnerf_hashgrid_encode(const rtxnerf_hashgrid* enc, const float* d_samples, int sample_stride,
                                       const long long* d_count, long long count_scale, long long n_max,
                                       uint16_t* d_encoded, void* stream)
{
    if (!enc || !d_samples || !d_encoded || sample_stride < 3 || n_max < 0) return RTXNERF_EINVAL;
    Levels lv;
    int rc = levels_from(enc, &lv);
    if (rc) return rc;
    if (n_max == 0) return RTXNERF_OK;
    long long blocks = (n_max + 255) / 256;
    if (blocks > RTX_NUM_SMS * 32) blocks = RTX_NUM_SMS * 32;
    hashgrid_encode_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(
        lv, reinterpret_cast<const __half2*>(enc->table), d_samples, sample_stride, d_count, count_scale, n_max,
        reinterpret_cast<__half2*>(d_encoded));
    return rtx_launch_status();
}

extern "C" int rtxnerf_network_forward(const rtxnerf_hashgrid* enc, const rtxnerf_mlp* mlp, const void* d_input,
                                       int input_stride, const long long* d_count, long long count_scale,
                                       long long n_max, uint16_t* d_out16, float* d_radiance4, int impl, void* stream)
{
    if (!mlp || !mlp->weights || !d_input || n_max < 0 || (!d_out16 && !d_radiance4)) return RTXNERF_EINVAL;
    if (mlp->in_dim != NET_IN || mlp->width != NET_W || mlp->n_hidden != 2) return RTXNERF_ECONFIG;
    if (mlp->out_activation != 0 && mlp->out_activation != 1) return RTXNERF_ECONFIG;
    Levels lv = {};
    if (enc) {
        if (input_stride < 3) return RTXNERF_EINVAL;
        int rc = levels_from(enc, &lv);
        if (rc) return rc;
        if (enc->n_levels * 2 > NET_IN) return RTXNERF_ECONFIG;
    }
    if (n_max == 0) return RTXNERF_OK;
    const __half2* table = enc ? reinterpret_cast<const __half2*>(enc->table) : nullptr;
    cudaStream_t s = (cudaStream_t)stream;
    if (impl == RTXNERF_IMPL_SIMT) {
        long long blocks = (n_max + 127) / 128;
        if (blocks > RTX_NUM_SMS * 16) blocks = RTX_NUM_SMS * 16;
        network_forward_simt_kernel<<<(unsigned)blocks, 128, 0, s>>>(lv, table, mlp->weights, d_input, input_stride,
                                                                   enc ? 1 : 0, mlp->out_activation, d_count,
                                                                   count_scale, n_max, d_out16, d_radiance4);
        return rtx_launch_status();
    }
    if (impl != RTXNERF_IMPL_TCGEN05) return RTXNERF_EINVAL;
    static RtxPerDeviceOnce attr_enc, attr_pre;
    RTX_CUDA_TRY((cudaError_t)rtx_set_max_smem(attr_enc, network_forward_tc_kernel<true>, NetSmem::BYTES));
    RTX_CUDA_TRY((cudaError_t)rtx_set_max_smem(attr_pre, network_forward_tc_kernel<false>, NetSmem::BYTES));
    long long tiles = (n_max + NET_TILE - 1) / NET_TILE;
    const long long resident = (long long)RTX_NUM_SMS * NET_CTAS_PER_SM;
    long long blocks = tiles < resident ? tiles : resident;
    if (enc)
        network_forward_tc_kernel<true><<<(unsigned)blocks, NET_TILE, NetSmem::BYTES, s>>>(
            lv, table, mlp->weights, d_input, input_stride, mlp->out_activation, d_count, count_scale, n_max, d_out16,
            d_radiance4);
    else
        network_forward_tc_kernel<false><<<(unsigned)blocks, NET_TILE, NetSmem::BYTES, s>>>(
            lv, table, mlp->weights, d_input, input_stride, mlp->out_activation, d_count, count_scale, n_max, d_out16,
            d_radiance4);
    return rtx_launch_status();
}
