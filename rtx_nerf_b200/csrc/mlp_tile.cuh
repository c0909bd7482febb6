// mlp_tile.cuh — the 128-sample MLP tile on tcgen05: shared-memory operand layout, layer
// issue and the TMEM epilogues, shared by the flat and the ray-coherent network kernels.
#pragma once
#include "common.cuh"
#include "tcgen05.cuh"

#define NET_TILE 128          // samples per tile = UMMA M
#define NET_IN 32
#define NET_W 64
#define NET_OUT 16
#define NET_TMEM_COLS 64      // one 128 x 64 fp32 accumulator, reused by all three layers (each layer's
                              // epilogue has drained it before the next MMA is issued)

// ---------------------------------------------------------------------------------------
// UMMA operand tiles in shared memory: element (row r, col k) of a [rows x K] fp16 tile at
//   (r>>3)*SBO + (k>>3)*128 + (r&7)*16 + (k&7)*2,   SBO = (K/8)*128, LBO = 128
// ---------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t tile_chunk_off(int r, int chunk, int K)
{
    return (uint32_t)((r >> 3) * (K * 16) + chunk * 128 + (r & 7) * 16);
}

// copy a row-major [rows][K] fp16 weight matrix from global into the UMMA layout
__device__ __forceinline__ void load_weight_tile(uint8_t* smem_tile, const uint16_t* __restrict__ w, int rows, int K,
                                                 int tid, int nthreads)
{
    const int chunks = K >> 3;
    for (int i = tid; i < rows * chunks; i += nthreads) {
        const int r = i / chunks, c = i - r * chunks;
        const uint4 v = *reinterpret_cast<const uint4*>(w + (size_t)r * K + c * 8);
        *reinterpret_cast<uint4*>(smem_tile + tile_chunk_off(r, c, K)) = v;
    }
}

// one layer: D[128 x N] = A[128 x K] * W[N x K]^T, issued by a single thread
template <int N, int K>
__device__ __forceinline__ void issue_layer(uint32_t d_tmem, uint32_t a_saddr, uint32_t w_saddr, uint64_t* bar)
{
    constexpr uint32_t idesc = tc::idesc_f16_f32(NET_TILE, N);
    constexpr uint32_t sbo = (K / 8) * 128;
#pragma unroll
    for (int ks = 0; ks < K / 16; ks++) {
        const uint64_t ad = tc::smem_desc(a_saddr + ks * 256, 128, sbo);
        const uint64_t bd = tc::smem_desc(w_saddr + ks * 256, 128, sbo);
        tc::mma_f16_ss(d_tmem, ad, bd, idesc, ks > 0 ? 1u : 0u);
    }
    tc::mma_commit(bar);
}

// two fp32 -> ReLU -> one packed fp16x2 in ONE instruction (F2FP.RELU): rounding is monotonic and
// keeps 0, so relu-then-round == round-then-relu bit for bit
__device__ __forceinline__ uint32_t relu_pack_f16x2(float lo, float hi)
{
    uint32_t d;
    asm("cvt.rn.relu.f16x2.f32 %0, %1, %2;" : "=r"(d) : "f"(hi), "f"(lo));
    return d;
}

// hidden-layer epilogue: TMEM row -> ReLU -> fp16 -> this thread's row of the next A tile.
// Both 16-column loads of a half row are issued before the single wait.
__device__ __forceinline__ void relu_row_to_smem(uint32_t taddr, uint8_t* a_next, int row)
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
#pragma unroll
        for (int c = 0; c < 4; c++)
            *reinterpret_cast<uint4*>(a_next + tile_chunk_off(row, 4 * h + c, NET_W)) =
                make_uint4(pk[4 * c], pk[4 * c + 1], pk[4 * c + 2], pk[4 * c + 3]);
    }
}

// output-layer epilogue: TMEM row (16 fp32) -> optional sigmoid -> fp16 -> out16[s][16] and/or
// radiance4[s] = float(half(.)) of the first four (main.cu:723-728, the intended gather of Q3)
__device__ __forceinline__ void output_row_load(uint32_t taddr, uint32_t (&r)[16])
{
    tc::tmem_ld_x16(taddr, r);
    tc::tmem_ld_wait();
}
// (r,g,b,sigma) of a row exactly as output_row_store writes them to radiance4: activation,
// one rounding to fp16, back to fp32
__device__ __forceinline__ float4 output_row_rgba(const uint32_t (&r)[16], int out_activation)
{
    float v[4];
#pragma unroll
    for (int j = 0; j < 2; j++) {
        float a = __uint_as_float(r[2 * j]), b = __uint_as_float(r[2 * j + 1]);
        if (out_activation == 1) {
            a = __fdividef(1.0f, 1.0f + __expf(-a));
            b = __fdividef(1.0f, 1.0f + __expf(-b));
        }
        const float2 back = __half22float2(__floats2half2_rn(a, b));
        v[2 * j] = back.x;
        v[2 * j + 1] = back.y;
    }
    return make_float4(v[0], v[1], v[2], v[3]);
}
__device__ __forceinline__ void output_row_store(const uint32_t (&r)[16], int out_activation, long long s,
                                                 uint16_t* __restrict__ out16, float* __restrict__ radiance4)
{
    uint32_t pk[8];
    float first4[4];
    const int n_pairs = out16 ? 8 : 2;     // only (r,g,b,sigma) are needed without out16
#pragma unroll
    for (int j = 0; j < 8; j++) {
        if (j >= n_pairs) break;
        float a = __uint_as_float(r[2 * j]), b = __uint_as_float(r[2 * j + 1]);
        if (out_activation == 1) {   // fast exp / divide: ~2 ulp in fp32, invisible after the fp16 rounding
            a = __fdividef(1.0f, 1.0f + __expf(-a));
            b = __fdividef(1.0f, 1.0f + __expf(-b));
        }
        const __half2 h = __floats2half2_rn(a, b);
        pk[j] = *reinterpret_cast<const uint32_t*>(&h);
        if (j < 2) {
            const float2 back = __half22float2(h);
            first4[2 * j] = back.x;
            first4[2 * j + 1] = back.y;
        }
    }
    if (out16) {
        uint4* o = reinterpret_cast<uint4*>(out16 + s * NET_OUT);
        o[0] = make_uint4(pk[0], pk[1], pk[2], pk[3]);
        o[1] = make_uint4(pk[4], pk[5], pk[6], pk[7]);
    }
    if (radiance4)
        st_stream_f4(reinterpret_cast<float4*>(radiance4) + s, make_float4(first4[0], first4[1], first4[2], first4[3]));
}
