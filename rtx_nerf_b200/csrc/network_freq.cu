// network_freq.cu — the reference's LITERAL network on tcgen05: Composite Frequency encoding +
// FullyFusedMLP 8 x 128, ReLU, sigmoid output, no bias, fp16 (config at main.cu:47-68, created at
// main.cu:325, forward at main.cu:715-728).  SURVEY §8 row f2.  tiny-cuda-nn is absent from the
// archive (PARITY UNPINNED): the published definitions are restated, see oracle/rtxnerf_oracle.c
// orc_frequency_encode / orc_mlp_forward.
//
// 262 144 FLOP per sample against 20 B in and 8..32 B out: this one IS tensor-bound, and the
// kernel is built so that the tensor pipe never waits for shared memory:
//   * activations never leave the tensor core's own memory.  A layer's accumulators
//     (128 x 128 fp32, 128 TMEM columns) are read with tcgen05.ld, ReLU'd and rounded to fp16
//     in registers, and written back with tcgen05.st as the A OPERAND of the next layer
//     (64 TMEM columns, A-from-TMEM form of tcgen05.mma).  With A in shared memory an
//     M128 N128 K16 MMA reads 8 KB per 64 cycles — the whole 128 B/clk of the SM's shared
//     memory — and the epilogue's stores would stall it; with A in TMEM shared memory only
//     feeds B (the weights).
//   * two tiles of 128 samples are in flight per SM (2 x (128 accumulator + 64 hidden-operand +
//     64 encoded-operand) = all 512 TMEM columns): while the tensor core multiplies tile 0 by
//     W_l, four warps run tile 1's epilogue, and eight more warps encode the NEXT pair of tiles
//     straight into TMEM, so the MUFU-bound encoding (108 sin/cos per row) is off the layer chain.
//   * the weights (256 KB, more than shared memory) are pre-packed once into the K-major
//     SWIZZLE_128B operand layout and streamed from L2 layer by layer through a 4-slot ring
//     with cp.async.bulk + mbarrier transaction counts; one load serves both tiles.
//   * one tcgen05.commit per layer-tile and nothing else on the issuing thread: a commit costs
//     it ~200 cycles, so ring slots are released by a plain mbarrier arrive from the epilogue.
// Measured (B200, 40.96 M rows): see DESIGN.md 4.7 / profiles/r01_e_freqnet_*.
#include "common.cuh"
#include "mlp_tile.cuh"

#define FN_XYZ_FREQS 10
#define FN_DIR_FREQS 12
#define FN_IN 112            // 3*10*2 + 2*12*2 = 108, padded with ones to a multiple of 16
#define FN_W 128
#define FN_HIDDEN 8
#define FN_LAYERS 9          // 112->128, 7 x 128->128, 128->16
#define FN_RING 4
#define FN_SLOT_BYTES (FN_W * FN_W * 2)
// Role layout (measured at 40.96 M rows, best of 10): 4 epilogue warps per slot + 4 encoder warps per slot
// 7.94 ms; 8 + 4 shared encoders 8.29 ms; 8 + 8 (72 registers, spills) 8.40 ms.
#ifndef FN_EPI_WARPS
#define FN_EPI_WARPS 4                          // epilogue warps per tile slot: 4, or 8 = two per TMEM lane quarter
#endif
#ifndef FN_ENC_WARPS
#define FN_ENC_WARPS 8                          // encoder warps: 8 (4 per slot) or 4 (both slots, alternately)
#endif
#define FN_EPI_COLS (FN_W / (FN_EPI_WARPS / 4)) // accumulator columns one epilogue warp converts
#define FN_WARP_ENC (2 * FN_EPI_WARPS)
#define FN_WARP_MMA (FN_WARP_ENC + FN_ENC_WARPS)
#define FN_WARP_LOAD (FN_WARP_MMA + 1)
#define FN_THREADS ((FN_WARP_LOAD + 1) * 32)
#define FN_PARAMS (FN_W * FN_IN + (FN_HIDDEN - 1) * FN_W * FN_W + NET_OUT * FN_W)     // row-major fp16 values
#define FN_PACKED_BYTES (FN_HIDDEN * FN_W * FN_W * 2 + NET_OUT * FN_W * 2)           // layer 0 padded to K = 128

__host__ __device__ constexpr int fn_layer_k(int l) { return l == 0 ? FN_IN : FN_W; }     // K of the row-major source
__host__ __device__ constexpr int fn_layer_n(int l) { return l == FN_LAYERS - 1 ? NET_OUT : FN_W; }
__host__ __device__ constexpr uint32_t fn_layer_bytes(int l) { return (uint32_t)(fn_layer_n(l) * FN_W * 2); }
__host__ __device__ constexpr uint32_t fn_layer_offset(int l) { return (uint32_t)(l * FN_W * FN_W * 2); }
__host__ __device__ constexpr uint32_t fn_layer_src(int l)      // first fp16 of layer l in the row-major weights
{
    return l == 0 ? 0u : (uint32_t)(FN_W * FN_IN + (l - 1) * FN_W * FN_W);
}

// row-major [out][in] per layer -> the K-major SWIZZLE_128B operand layout, layers concatenated
// (a ring slot is filled by ONE bulk copy and read by the MMA as is):
//   element (r, k) of an [N][128] layer at  (k/64)*(N*128) + (r/8)*1024 + (r%8)*128 + (((k%64)/8) ^ (r%8))*16 + (k%8)*2
// i.e. two K-blocks of 64 (one 128-byte row each), 8-row swizzle atoms of 1 KB.  Layer 0's 112
// inputs are zero-padded to 128.  (The no-swizzle core-matrix layout of mlp_tile.cuh ran the
// M128 N128 K16 MMAs at ~40 % of their rate here.)
__global__ void freqnet_pack_kernel(const uint16_t* __restrict__ w, uint8_t* __restrict__ packed)
{
    const int total_chunks = FN_PACKED_BYTES / 16;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total_chunks; i += gridDim.x * blockDim.x) {
        const int l = min(i / (FN_W * FN_W * 2 / 16), FN_LAYERS - 1);
        const int N = fn_layer_n(l), K = fn_layer_k(l);
        const int local = i - (int)(fn_layer_offset(l) / 16);      // chunk index (r, c) in [N][16 chunks of 8]
        const int r = local >> 4, c = local & 15;
        uint4 v = make_uint4(0u, 0u, 0u, 0u);
        if (c * 8 < K) v = *reinterpret_cast<const uint4*>(w + fn_layer_src(l) + (size_t)r * K + c * 8);
        const uint32_t off = (uint32_t)((c >> 3) * (N * 128) + (r >> 3) * 1024 + (r & 7) * 128 + (((c & 7) ^ (r & 7)) << 4));
        *reinterpret_cast<uint4*>(packed + fn_layer_offset(l) + off) = v;
    }
}

// sin(pi*y), cos(pi*y) packed as fp16x2 {lo = sin, hi = cos}.  y - 2*rint(y/2) is exact in
// fp32, so the hardware sine sees an argument in [-pi, pi] whatever the octave.
__device__ __forceinline__ uint32_t sincospi_f16x2(float y)
{
    const float r = rn_fma(-2.0f, rintf(rn_mul(y, 0.5f)), y);
    const float a = rn_mul(r, 3.14159265358979323846f);
    const __half2 h = __floats2half2_rn(__sinf(a), __cosf(a));
    return *reinterpret_cast<const uint32_t*>(&h);
}

// the 56 packed columns (112 features) of one sample row, main.cu:47-61 order:
// x: f0..f9 (sin, cos), y, z, theta: f0..f11, phi, then ones
template <class Sink>
__device__ __forceinline__ void frequency_row(const float* __restrict__ p, bool valid, Sink sink)
{
    uint32_t col[16];
    int nc = 0, base = 0;
#pragma unroll
    for (int d = 0; d < 5; d++) {
        float y = valid ? __ldg(p + d) : 0.0f;
        const int F = d < 3 ? FN_XYZ_FREQS : FN_DIR_FREQS;
#pragma unroll
        for (int k = 0; k < F; k++) {
            col[nc++] = valid ? sincospi_f16x2(y) : 0u;
            y = rn_add(y, y);
            if (nc == 16) { sink(base, col); base += 16; nc = 0; }
        }
    }
    // 54 columns so far: 6 in col[]; two columns of ones complete the 112 features, eight
    // columns of zeros pad the row to K = 128 for the tensor core (the weights there are zero)
    col[6] = valid ? 0x3c003c00u : 0u;
    col[7] = col[6];
#pragma unroll
    for (int i = 8; i < 16; i++) col[i] = 0u;
    sink.tail(base, col);
}

struct TmemRowSink {
    uint32_t taddr;   // this thread's lane row, first A column
    __device__ __forceinline__ void operator()(int base, const uint32_t (&c)[16]) const { tc::tmem_st_x16(taddr + base, c); }
    __device__ __forceinline__ void tail(int base, const uint32_t (&c)[16]) const { tc::tmem_st_x16(taddr + base, c); }
};
struct GlobalRowSink {
    uint32_t* row;    // [56] packed columns of this sample, or nullptr
    __device__ __forceinline__ void operator()(int base, const uint32_t (&c)[16]) const
    {
#pragma unroll
        for (int i = 0; i < 4; i++)
            reinterpret_cast<uint4*>(row + base)[i] = make_uint4(c[4 * i], c[4 * i + 1], c[4 * i + 2], c[4 * i + 3]);
    }
    __device__ __forceinline__ void tail(int base, const uint32_t (&c)[16]) const   // the 112 real features only
    {
#pragma unroll
        for (int i = 0; i < 2; i++)
            reinterpret_cast<uint4*>(row + base)[i] = make_uint4(c[4 * i], c[4 * i + 1], c[4 * i + 2], c[4 * i + 3]);
    }
};

// encoding alone (tests, and the pre-encoded path of callers that want the features)
__global__ void freqnet_encode_kernel(const float* __restrict__ samples, int stride, long long n,
                                      uint16_t* __restrict__ encoded)
{
    const long long s = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n) return;
    GlobalRowSink sink{reinterpret_cast<uint32_t*>(encoded + s * FN_IN)};
    frequency_row(samples + s * stride, true, sink);
}

#ifdef FN_TRACE
// development trace: clock64 stamps of CTA 0 (role 0/1 = epilogue of slot 0/1, role 2 = MMA thread)
__device__ long long fn_trace[3 * 1024];
__device__ __forceinline__ void fn_mark(int role, int& idx)
{
    if (blockIdx.x == 0 && idx < 1024) fn_trace[role * 1024 + idx] = clock64();
    idx++;
}
extern "C" int rtxnerf_debug_freqnet_trace(long long* host_out)
{
    return (int)cudaMemcpyFromSymbol(host_out, fn_trace, sizeof(fn_trace));
}
#else
__device__ __forceinline__ void fn_mark(int, int&) {}
#endif

struct FnSmem {
    static constexpr int RING = 0;                                   // FN_RING slots of 32 KB, 1 KB-aligned
    static constexpr int BAR = RING + FN_RING * FN_SLOT_BYTES;       // see FnBars
    static constexpr int TMEM_SLOT = BAR + 8 * (2 * FN_RING + 10);
    static constexpr int BYTES = TMEM_SLOT + 16 + 1024;             // + slack for the 1 KB alignment
};
// TMEM columns of tile slot s (base s*256): accumulators, hidden-layer A operand, encoded A operand
#define FN_COL_D 0
#define FN_COL_AH 128
#define FN_COL_AE 192

// hidden-layer epilogue of one row: 128 fp32 accumulator columns -> ReLU -> fp16 -> 64 A columns
__device__ __forceinline__ void fn_hidden_epilogue(uint32_t t_d, uint32_t t_a, int h_begin)
{
#pragma unroll
    for (int h = h_begin; h < h_begin + FN_EPI_COLS / 32; h++) {
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
}

// Roles (18 warps by default): 0-3 / 4-7 epilogue of tile slot 0 / 1, 8-11 / 12-15 encoder of slot
// 0 / 1 (warp % 4 = TMEM lane quarter everywhere), 16 MMA issue, 17 weight loader.
// Hand-offs (mbarriers): enc_ready[s] encoder -> MMA (layer-0 operand written), enc_free[s]
// epilogue -> encoder (layer-0 MMA of the previous tile has completed: its operand may be
// overwritten), a_ready[s] epilogue -> MMA (next hidden operand written), d_ready[s] MMA ->
// epilogue (tcgen05.commit: accumulators complete), w_full[k] loader -> MMA (bulk copy landed),
// w_empty[k] slot-1 epilogue -> loader (the layer's MMAs of BOTH tiles have completed: the tensor
// pipe runs in order, so tile 1's completion covers tile 0), o_free[s] epilogue -> MMA (the output
// accumulators of the previous tile have been read: layer 0 of the next tile overwrites them, and
// with the encoder running ahead nothing else orders the two).  The only tcgen05.commit per
// layer-tile is d_ready: a commit stalls the issuing thread for ~200 cycles.
__global__ void __launch_bounds__(FN_THREADS, 1)
freqnet_forward_kernel(const uint8_t* __restrict__ packed, const float* __restrict__ samples, int stride,
                       int out_activation, const long long* __restrict__ d_count, long long count_scale,
                       long long n_max, uint16_t* __restrict__ out16, float* __restrict__ radiance4)
{
    extern __shared__ __align__(128) uint8_t smem_raw[];
    uint8_t* smem = smem_raw + ((1024u - (tc::smem_u32(smem_raw) & 1023u)) & 1023u);   // swizzle atoms are 1 KB
    uint64_t* w_full = reinterpret_cast<uint64_t*>(smem + FnSmem::BAR);
    uint64_t* w_empty = w_full + FN_RING;
    uint64_t* a_ready = w_empty + FN_RING;
    uint64_t* d_ready = a_ready + 2;
    uint64_t* enc_ready = d_ready + 2;
    uint64_t* enc_free = enc_ready + 2;
    uint64_t* o_free = enc_free + 2;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + FnSmem::TMEM_SLOT);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;

    if (tid == 0) {
        for (int i = 0; i < FN_RING; i++) {
            tc::mbar_init(w_full + i, 1);
            tc::mbar_init(w_empty + i, 1);
        }
        for (int i = 0; i < 2; i++) {
            tc::mbar_init(a_ready + i, FN_EPI_WARPS * 32);
            tc::mbar_init(d_ready + i, 1);
            tc::mbar_init(enc_ready + i, NET_TILE);
            tc::mbar_init(enc_free + i, 1);
            tc::mbar_init(o_free + i, NET_TILE);
        }
        tc::fence_mbar_init();
    }
    if (warp == 0) tc::tmem_alloc<512>(tmem_slot);
    tc::fence_before_thread_sync();
    __syncthreads();
    tc::fence_after_thread_sync();
    const uint32_t tmem_base = *tmem_slot;

    long long n = n_max;
    if (d_count) {
        const long long c = *d_count * count_scale;
        n = c < n_max ? c : n_max;
    }
    const long long n_tiles = (n + NET_TILE - 1) / NET_TILE;
    const long long n_pairs = (n_tiles + 1) / 2;

    if (warp < FN_WARP_ENC) {
        // ------------------------------------------------ epilogue group of tile slot s: warp & 3 = TMEM
        // lane quarter, half = which 64 of the 128 accumulator columns this warp converts
        const int s = warp / FN_EPI_WARPS, wk = warp & 3, half = FN_EPI_WARPS == 8 ? (warp >> 2) & 1 : 0;
        const uint32_t t_d = tmem_base + (uint32_t)(s * 256 + FN_COL_D) + ((uint32_t)(wk * 32) << 16);
        const uint32_t t_a = t_d + FN_COL_AH;
        const int row_in_tile = wk * 32 + lane;
        const bool leader = wk == 0 && half == 0 && lane == 0;
        uint32_t d_par = 0, cnt = 0;
        int tr = 0;
        for (long long pair = blockIdx.x; pair < n_pairs; pair += gridDim.x) {
            const long long row = (pair * 2 + s) * NET_TILE + row_in_tile;
            for (int l = 0; l < FN_HIDDEN; l++, cnt++) {
                tc::mbar_wait(d_ready + s, d_par);
                d_par ^= 1;
                if (leader) {
                    fn_mark(s, tr);                               // accumulators seen
                    if (l == 0) tc::mbar_arrive(enc_free + s);    // the encoded operand has been consumed
                    if (s == 1) tc::mbar_arrive(w_empty + cnt % FN_RING);
                }
                tc::fence_after_thread_sync();
                fn_hidden_epilogue(t_d, t_a, half * (FN_EPI_COLS / 32));
                tc::tmem_st_wait();
                tc::fence_before_thread_sync();
                tc::mbar_arrive(a_ready + s);
                if (leader) fn_mark(s, tr);                       // next operand published
            }
            tc::mbar_wait(d_ready + s, d_par);
            d_par ^= 1;
            if (leader && s == 1) tc::mbar_arrive(w_empty + cnt % FN_RING);
            cnt++;
            tc::fence_after_thread_sync();
            if (half == 0) {                                      // 16 output columns: one warp per lane quarter
                uint32_t r[16];
                output_row_load(t_d, r);
                tc::fence_before_thread_sync();
                tc::mbar_arrive(o_free + s);                      // the next tile's layer 0 may overwrite D
                if (row < n) output_row_store(r, out_activation, row, out16, radiance4);
            }
        }
    } else if (warp < FN_WARP_MMA) {
        // ------------------------------------------------ encoders: four warps write the layer-0 operand
        // of slot 0, then of slot 1 (one tile pair needs them for ~1/4 of its 9-layer period)
        const int wk = warp & 3;
        const int row_in_tile = wk * 32 + lane;
        const int s_begin = FN_ENC_WARPS == 8 ? (warp - FN_WARP_ENC) >> 2 : 0;
        const int s_end = FN_ENC_WARPS == 8 ? s_begin + 1 : 2;
        uint32_t f_par = 0;
        bool first = true;
        for (long long pair = blockIdx.x; pair < n_pairs; pair += gridDim.x) {
#pragma unroll 1
            for (int s = s_begin; s < s_end; s++) {
                const uint32_t t_e = tmem_base + (uint32_t)(s * 256 + FN_COL_AE) + ((uint32_t)(wk * 32) << 16);
                const long long row = (pair * 2 + s) * NET_TILE + row_in_tile;
                const bool valid = row < n;
                if (!first) {
                    tc::mbar_wait(enc_free + s, f_par);
                    tc::fence_after_thread_sync();
                }
                TmemRowSink sink{t_e};
                frequency_row(samples + (valid ? row : 0) * stride, valid, sink);
                tc::tmem_st_wait();
                tc::fence_before_thread_sync();
                tc::mbar_arrive(enc_ready + s);
            }
            if (!first) f_par ^= 1;
            first = false;
        }
    } else if (warp == FN_WARP_MMA) {
        // ------------------------------------------------ MMA issue (one elected lane: tcgen05.cuh elect_one)
        if (tc::elect_one()) {
            uint32_t cnt = 0, a_par0 = 0, a_par1 = 0, e_par = 0, o_par = 0;
            bool first_pair = true;
            int tr = 0;
            const uint32_t ring = tc::smem_u32(smem + FnSmem::RING);
            for (long long pair = blockIdx.x; pair < n_pairs; pair += gridDim.x) {
#pragma unroll 1
                for (int l = 0; l < FN_LAYERS; l++, cnt++) {
                    const uint32_t slot = cnt % FN_RING;
                    tc::mbar_wait(w_full + slot, (cnt / FN_RING) & 1u);
                    const uint32_t idesc = l == FN_LAYERS - 1 ? tc::idesc_f16_f32(NET_TILE, NET_OUT)
                                                               : tc::idesc_f16_f32(NET_TILE, FN_W);
                    const uint32_t kb_stride = (uint32_t)fn_layer_n(l) * 128u;     // second K-block of 64
                    const uint32_t w_addr = ring + slot * FN_SLOT_BYTES;
#pragma unroll
                    for (int s = 0; s < 2; s++) {
                        if (l == 0) {
                            tc::mbar_wait(enc_ready + s, e_par);
                            // the previous tile's output accumulators (D columns 0-15) have been read
                            if (!first_pair) tc::mbar_wait(o_free + s, o_par);
                        } else {
                            tc::mbar_wait(a_ready + s, s ? a_par1 : a_par0);
                            if (s) a_par1 ^= 1; else a_par0 ^= 1;
                        }
                        tc::fence_after_thread_sync();
                        fn_mark(2, tr);                       // per (l, s): operand seen
                        const uint32_t t_d = tmem_base + (uint32_t)(s * 256 + FN_COL_D);
                        const uint32_t t_a = tmem_base + (uint32_t)(s * 256 + (l == 0 ? FN_COL_AE : FN_COL_AH));
#pragma unroll
                        for (int ks = 0; ks < FN_W / 16; ks++)
                            tc::mma_f16_ts(t_d, t_a + ks * 8,
                                           tc::smem_desc_sw128(w_addr + (ks >> 2) * kb_stride + (ks & 3) * 32), idesc,
                                           ks > 0 ? 1u : 0u);
                        tc::mma_commit(d_ready + s);
                        fn_mark(2, tr);                       // per (l, s): MMAs issued + committed
                    }
                    if (l == 0) {
                        e_par ^= 1;
                        if (!first_pair) o_par ^= 1;
                        first_pair = false;
                    }
                }
            }
        }
    } else {
        // ------------------------------------------------ weight loader
        if (lane == 0) {
            uint32_t cnt = 0;
            for (long long pair = blockIdx.x; pair < n_pairs; pair += gridDim.x) {
#pragma unroll 1
                for (int l = 0; l < FN_LAYERS; l++, cnt++) {
                    const uint32_t slot = cnt % FN_RING, use = cnt / FN_RING;
                    if (use > 0) tc::mbar_wait(w_empty + slot, (use - 1) & 1u);
                    const uint32_t bytes = fn_layer_bytes(l);
                    tc::mbar_arrive_expect_tx(w_full + slot, bytes);
                    tc::bulk_g2s(smem + FnSmem::RING + slot * FN_SLOT_BYTES, packed + fn_layer_offset(l), bytes,
                                 w_full + slot);
                }
            }
        }
    }
    tc::fence_before_thread_sync();
    __syncthreads();
    tc::fence_after_thread_sync();
    if (warp == 0) tc::tmem_dealloc<512>(tmem_base);
}

// ------------------------------------------------------------------------------------ host ABI
extern "C" size_t rtxnerf_freqnet_param_count(void) { return (size_t)FN_PARAMS; }
extern "C" size_t rtxnerf_freqnet_packed_count(void) { return (size_t)FN_PACKED_BYTES / 2; }

extern "C" int rtxnerf_freqnet_pack(const uint16_t* d_weights, uint16_t* d_packed, void* stream)
{
    if (!d_weights || !d_packed) return RTXNERF_EINVAL;
    if (((uintptr_t)d_weights & 15u) || ((uintptr_t)d_packed & 15u)) return RTXNERF_EINVAL;
    freqnet_pack_kernel<<<64, 256, 0, (cudaStream_t)stream>>>(d_weights, reinterpret_cast<uint8_t*>(d_packed));
    return rtx_launch_status();
}

static int freqnet_check(const rtxnerf_freqnet* net)
{
    if (!net || !net->weights_packed) return RTXNERF_EINVAL;
    if (net->n_freq_xyz != FN_XYZ_FREQS || net->n_freq_dir != FN_DIR_FREQS || net->width != FN_W ||
        net->n_hidden != FN_HIDDEN)
        return RTXNERF_ECONFIG;
    if (net->out_activation != 0 && net->out_activation != 1) return RTXNERF_ECONFIG;
    if ((uintptr_t)net->weights_packed & 15u) return RTXNERF_EINVAL;
    return RTXNERF_OK;
}

extern "C" int rtxnerf_freqnet_encode(const float* d_samples, int sample_stride, long long n, uint16_t* d_encoded,
                                      void* stream)
{
    if (!d_samples || !d_encoded || n < 0 || sample_stride < 5) return RTXNERF_EINVAL;
    if ((uintptr_t)d_encoded & 15u) return RTXNERF_EINVAL;
    if (n == 0) return RTXNERF_OK;
    freqnet_encode_kernel<<<(unsigned)((n + 127) / 128), 128, 0, (cudaStream_t)stream>>>(d_samples, sample_stride, n,
                                                                                         d_encoded);
    return rtx_launch_status();
}

extern "C" int rtxnerf_freqnet_forward(const rtxnerf_freqnet* net, const float* d_samples, int sample_stride,
                                       const long long* d_count, long long count_scale, long long n_max,
                                       uint16_t* d_out16, float* d_radiance4, void* stream)
{
    int rc = freqnet_check(net);
    if (rc) return rc;
    if (!d_samples || (!d_out16 && !d_radiance4) || n_max < 0 || sample_stride < 5 || count_scale < 1)
        return RTXNERF_EINVAL;
    if (n_max == 0) return RTXNERF_OK;
    static RtxPerDeviceOnce attr;
    RTX_CUDA_TRY((cudaError_t)rtx_set_max_smem(attr, freqnet_forward_kernel, FnSmem::BYTES));
    const long long pairs = ((n_max + NET_TILE - 1) / NET_TILE + 1) / 2;
    const long long blocks = pairs < RTX_NUM_SMS ? pairs : RTX_NUM_SMS;   // one CTA per SM owns all 512 TMEM columns
    freqnet_forward_kernel<<<(unsigned)blocks, FN_THREADS, FnSmem::BYTES, (cudaStream_t)stream>>>(
        reinterpret_cast<const uint8_t*>(net->weights_packed), d_samples, sample_stride, net->out_activation, d_count,
        count_scale, n_max, d_out16, d_radiance4);
    return rtx_launch_status();
}

// =========================================================================================
// Backward of the reference's literal network (main.cu:781 network->backward): the weight
// gradients of the 8 x 128 MLP over the Frequency encoding (the encoding has no parameters, so nothing
// flows below layer 0).  tiny-cuda-nn is absent from the archive: PARITY UNPINNED, the algorithm is
// restated (oracle: orc_freqnet_backward; independent autograd pin: tests/test_oracle_autograd.py).
//   dOut_j = g_j * sigmoid'(out_j);  dZ_{l-1} = (dZ_l . W_l) masked by ReLU'(X_l);  dW_l += dZ_l^T . X_l
// The forward is recomputed per 128-sample tile (keeping the activations would cost 2 KB per sample).
// The nine weight-gradient accumulators are 9 x 128 fp32 TMEM columns — more than the 512 an SM has next
// to the 128 of the working accumulator — so the kernel is launched THREE times, each launch recomputing
// the forward and the dgrad chain and accumulating the gradients of three layers in TMEM over all of its
// tiles (one flush of fp32 atomics per CTA at the end).  A launch stops its dgrad chain at its lowest kept layer:
// 27 forward + 15 dgrad + 9 wgrad = 51 layer products per tile instead of the minimal 27.
// Operand tiles live in shared memory in the no-swizzle core-matrix layout, which the tensor cores read in
// both majors (tcgen05.cuh): X_l [sample][feature] is the K-major A of the forward layer and the MN-major
// B of dW_l; dZ_l is the K-major A of the dgrad and the MN-major A of dW_l; W_l [out][in] is the K-major B
// of the forward and the MN-major B of the dgrad.  Weights stream from L2 into a two-slot ring by plain
// cp.async copies issued under the running MMA.  One tile per CTA, serial phases; the measured rate is in
// DESIGN.md.
// =========================================================================================
#define FB_THREADS (2 * NET_TILE)          // two threads per sample row: each owns 64 of the 128 columns
#define FB_TILE_BYTES (NET_TILE * FN_W * 2)                 // 32 KB: [128][128] fp16
struct FbSmem {
    static constexpr int WB = 0;                             // two weight slots
    static constexpr int XK = WB + 2 * FB_TILE_BYTES;        // inputs X_l of the three layers of this launch
    static constexpr int WORK = XK + 3 * FB_TILE_BYTES;      // current X_l, then current dZ_l
    static constexpr int DOUT = WORK + FB_TILE_BYTES;        // [128][16] dOut
    static constexpr int BAR = DOUT + NET_TILE * NET_OUT * 2;
    static constexpr int BYTES = BAR + 16;
};
#define FB_COL_D 0
#define FB_COL_G(i) (128 + 128 * (i))

// row-major [rows][K] fp16 layer -> core-matrix tile with 128 columns (columns >= K are zero).  16-byte
// cp.async copies: a chunk of eight K values of one row IS one row of a core matrix, so the layout change is
// only the destination address; no registers, nothing for a thread to wait on until fb_weights_landed().
// (Through registers the loads were 24 % of the kernel's stall samples and, with the compiler hoisting them
// above barriers, the reason it ran at 255 registers + spills; ncu, profiles/r02_cc_freqnet_backward_lines.txt.)
__device__ __forceinline__ void fb_load_weights(uint8_t* tile, const uint16_t* __restrict__ w, int l, int tid)
{
    const int rows = fn_layer_n(l), K = fn_layer_k(l);
    const uint16_t* src = w + fn_layer_src(l);
    const uint32_t dst = tc::smem_u32(tile);
#pragma unroll 4
    for (int i = tid; i < rows * 16; i += FB_THREADS) {
        const int r = i >> 4, c = i & 15;
        const bool real = c * 8 < K;             // padding columns: zero fill (source size 0)
        const uint16_t* g = real ? src + (size_t)r * K + c * 8 : src;
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst + tile_chunk_off(r, c, FN_W)), "l"(g),
                     "r"(real ? 16u : 0u)
                     : "memory");
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
}
__device__ __forceinline__ void fb_weights_landed() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }

struct SmemRowSink {       // frequency_row -> this thread's row of a [128][128] core-matrix tile
    uint8_t* tile;
    int row;
    __device__ __forceinline__ void operator()(int base, const uint32_t (&c)[16]) const
    {
#pragma unroll
        for (int i = 0; i < 4; i++)
            *reinterpret_cast<uint4*>(tile + tile_chunk_off(row, base / 4 + i, FN_W)) =
                make_uint4(c[4 * i], c[4 * i + 1], c[4 * i + 2], c[4 * i + 3]);
    }
    __device__ __forceinline__ void tail(int base, const uint32_t (&c)[16]) const { (*this)(base, c); }
};

// The three issue helpers are NOT inlined: inlined into the 17 unrolled phases, their ~250 64-bit matrix
// descriptors (all derived from a few shared-memory bases) were computed ahead and kept live by every thread
// of the kernel — 255 registers and 2.8 KB of spill — although only thread 0 issues.
// D[128 x N] = A[128 x 128] (K-major) . B[N x 128]^T (K-major)
template <int N>
__device__ __noinline__ void fb_mma_fwd(uint32_t d, uint32_t a, uint32_t w)
{
    constexpr uint32_t id = tc::idesc_f16_f32_major(NET_TILE, N, 0, 0);
    const uint64_t da = tc::smem_desc(a, 128, FN_W * 16), dw = tc::smem_desc(w, 128, FN_W * 16);
#pragma unroll
    for (int ks = 0; ks < FN_W / 16; ks++)
        tc::mma_f16_ss(d, tc::smem_desc_advance(da, ks * 256), tc::smem_desc_advance(dw, ks * 256), id, ks > 0 ? 1u : 0u);
}
// D[128 x 128] = dZ[128 x K_OUT] (K-major, tile of K_OUT columns) . W[K_OUT out][128 in] (MN-major)
template <int K_OUT>
__device__ __noinline__ void fb_mma_dgrad(uint32_t d, uint32_t dz, uint32_t w)
{
    constexpr uint32_t id = tc::idesc_f16_f32_major(NET_TILE, FN_W, 0, 1);
    const uint64_t dd = tc::smem_desc(dz, 128, K_OUT * 16), dw = tc::smem_desc(w, FN_W * 16, 128);
#pragma unroll
    for (int ks = 0; ks < K_OUT / 16; ks++)
        tc::mma_f16_ss(d, tc::smem_desc_advance(dd, ks * 256), tc::smem_desc_advance(dw, ks * 2 * (FN_W * 16)), id,
                       ks > 0 ? 1u : 0u);
}
// G[128 x N] (+)= X[128 samples][128]^T . Y[128 samples][N]   (both MN-major, K = samples)
template <int N>
__device__ __noinline__ void fb_mma_wgrad(uint32_t g, uint32_t x, uint32_t y, bool accumulate)
{
    constexpr uint32_t id = tc::idesc_f16_f32_major(NET_TILE, N, 1, 1);
    const uint64_t dx = tc::smem_desc(x, FN_W * 16, 128), dy = tc::smem_desc(y, N * 16, 128);
#pragma unroll
    for (int ks = 0; ks < NET_TILE / 16; ks++)
        tc::mma_f16_ss(g, tc::smem_desc_advance(dx, ks * 2 * (FN_W * 16)), tc::smem_desc_advance(dy, ks * 2 * (N * 16)), id,
                       (accumulate || ks > 0) ? 1u : 0u);
}

// forward epilogue of one half row (columns 64 * half ..): fp32 -> ReLU -> fp16 -> dst; mask = (activation > 0)
__device__ __forceinline__ void fb_relu_row(uint32_t t_row, uint8_t* dst, int row, int half, uint32_t (&mask)[2])
{
#pragma unroll
    for (int hh = 0; hh < FN_W / 64; hh++) {        // 32 columns per step: both loads are issued before the wait
        const int q = 4 * half + 2 * hh;
        uint32_t r[32];
        tc::tmem_ld_x16(t_row + q * 16, *reinterpret_cast<uint32_t(*)[16]>(r));
        tc::tmem_ld_x16(t_row + q * 16 + 16, *reinterpret_cast<uint32_t(*)[16]>(r + 16));
        tc::tmem_ld_wait();
        uint32_t pk[16], m = 0u;
#pragma unroll
        for (int j = 0; j < 16; j++) {
            pk[j] = relu_pack_f16x2(__uint_as_float(r[2 * j]), __uint_as_float(r[2 * j + 1]));
            m |= ((pk[j] & 0xffffu) ? 1u : 0u) << (2 * j);
            m |= ((pk[j] >> 16) ? 1u : 0u) << (2 * j + 1);
        }
        mask[hh] = m;
#pragma unroll
        for (int c = 0; c < 4; c++)
            *reinterpret_cast<uint4*>(dst + tile_chunk_off(row, 2 * q + c, FN_W)) =
                make_uint4(pk[4 * c], pk[4 * c + 1], pk[4 * c + 2], pk[4 * c + 3]);
    }
}
// dgrad epilogue of one half row: fp32 -> masked -> fp16 -> dst
__device__ __forceinline__ void fb_masked_row(uint32_t t_row, uint8_t* dst, int row, int half, const uint32_t (&mask)[2])
{
#pragma unroll
    for (int hh = 0; hh < FN_W / 64; hh++) {
        const int q = 4 * half + 2 * hh;
        uint32_t r[32];
        tc::tmem_ld_x16(t_row + q * 16, *reinterpret_cast<uint32_t(*)[16]>(r));
        tc::tmem_ld_x16(t_row + q * 16 + 16, *reinterpret_cast<uint32_t(*)[16]>(r + 16));
        tc::tmem_ld_wait();
        const uint32_t m = mask[hh];
        uint32_t pk[16];
#pragma unroll
        for (int j = 0; j < 16; j++) {
            const float a = ((m >> (2 * j)) & 1u) ? __uint_as_float(r[2 * j]) : 0.0f;
            const float b = ((m >> (2 * j + 1)) & 1u) ? __uint_as_float(r[2 * j + 1]) : 0.0f;
            const __half2 v = __floats2half2_rn(a, b);
            pk[j] = *reinterpret_cast<const uint32_t*>(&v);
        }
#pragma unroll
        for (int c = 0; c < 4; c++)
            *reinterpret_cast<uint4*>(dst + tile_chunk_off(row, 2 * q + c, FN_W)) =
                make_uint4(pk[4 * c], pk[4 * c + 1], pk[4 * c + 2], pk[4 * c + 3]);
    }
}

__global__ void __launch_bounds__(FB_THREADS, 1)
freqnet_backward_kernel(const uint16_t* __restrict__ weights, const float* __restrict__ samples, int stride,
                        int out_activation, const uint16_t* __restrict__ dl_dradiance,
                        const long long* __restrict__ d_count, long long count_scale, long long n_max, int l0,
                        float* __restrict__ grad_weights)
{
    extern __shared__ __align__(128) uint8_t smem[];
    uint64_t* bar = reinterpret_cast<uint64_t*>(smem + FbSmem::BAR);
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + FbSmem::BAR + 8);
    const int tid = threadIdx.x, warp = tid >> 5;
    const int row = tid & (NET_TILE - 1), half = tid >> 7;      // this thread's sample row and column half
    if (tid == 0) {
        tc::mbar_init(bar, 1);
        tc::fence_mbar_init();
    }
    if (warp == 0) tc::tmem_alloc<512>(tmem_slot);
    fb_load_weights(smem + FbSmem::WB, weights, 0, tid);
    fb_weights_landed();
    tc::fence_proxy_async_smem();
    tc::fence_before_thread_sync();
    __syncthreads();
    tc::fence_after_thread_sync();
    const uint32_t tm = *tmem_slot;
    const uint32_t t_row = tm + ((uint32_t)((warp & 3) * 32) << 16);     // TMEM lane quarter = warp % 4
    uint32_t phase = 0;
    bool have = false;      // the three gradient accumulators hold at least one tile

    // publish this phase's shared-memory tile, run the MMAs, do UNDER while they run, wait for them
    // (a macro, not a lambda: by-reference captures pushed the mask registers into local memory)
#define FB_RUN(ISSUE, UNDER)                        \
    do {                                            \
        fb_weights_landed();                        \
        tc::fence_proxy_async_smem();               \
        tc::fence_before_thread_sync();             \
        __syncthreads();                            \
        if (warp == 0 && tc::elect_one()) {         \
            tc::fence_after_thread_sync();          \
            ISSUE;                                  \
            tc::mma_commit(bar);                    \
        }                                           \
        UNDER;                                      \
        tc::mbar_wait(bar, phase);                  \
        phase ^= 1;                                 \
        tc::fence_after_thread_sync();              \
    } while (0)
#define FB_KEEP(l) ((l) >= l0 && (l) < l0 + 3)
#define FB_X(l) (FB_KEEP(l) ? sm + FbSmem::XK + ((l) - l0) * FB_TILE_BYTES : work)

    long long n = n_max;
    if (d_count) {
        const long long c = *d_count * count_scale;
        n = c < n_max ? c : n_max;
    }
    const long long n_tiles = (n + NET_TILE - 1) / NET_TILE;
    for (long long tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        // every operand address of the tile derives from `sm`, opaque per iteration: left alone, the compiler
        // hoists the ~250 loop-invariant 64-bit matrix descriptors of the 17 phases out of the tile loop and
        // spills them (255 registers, 2.8 KB of stack)
        uint8_t* sm = smem;
        asm volatile("" : "+l"(sm));
        uint8_t* wb0 = sm + FbSmem::WB;
        uint8_t* wb1 = sm + FbSmem::WB + FB_TILE_BYTES;
        uint8_t* work = sm + FbSmem::WORK;
        uint8_t* dout = sm + FbSmem::DOUT;
        const long long s = tile * NET_TILE + row;
        const bool valid = s < n;
        uint32_t mask[FN_HIDDEN][2];        // mask[l] = (X_{l+1} > 0), one bit per column of this thread's half row
        if (half == 0) frequency_row(samples + (valid ? s : 0) * stride, valid, SmemRowSink{FB_X(0), row});
        // ---- forward recompute: X_{l+1} = relu(X_l . W_l^T); W_{l+1} is loaded under the MMA of layer l
#pragma unroll
        for (int l = 0; l < FN_HIDDEN; l++) {
            uint8_t* xc = FB_X(l);
            uint8_t* xn = FB_X(l + 1);
            uint8_t* wc = (l & 1) ? wb1 : wb0;
            uint8_t* wn = (l & 1) ? wb0 : wb1;
            FB_RUN(fb_mma_fwd<FN_W>(tm + FB_COL_D, tc::smem_u32(xc), tc::smem_u32(wc)),
                   fb_load_weights(wn, weights, l + 1, tid));
            fb_relu_row(t_row + FB_COL_D, xn, row, half, mask[l]);
        }
        uint8_t* x8 = FB_X(FN_HIDDEN);
        FB_RUN(fb_mma_fwd<NET_OUT>(tm + FB_COL_D, tc::smem_u32(x8), tc::smem_u32(wb0)), (void)0);
        if (half == 0) {   // dOut = g * act'(out) on (r,g,b,sigma); the 12 padding outputs carry no gradient
            uint32_t r[16];
            output_row_load(t_row + FB_COL_D, r);
            float d[4] = {0.f, 0.f, 0.f, 0.f};
            if (valid) {
                const uint2 gb = __ldg(reinterpret_cast<const uint2*>(dl_dradiance + s * 4));
                const float2 g01 = __half22float2(*reinterpret_cast<const __half2*>(&gb.x));
                const float2 g23 = __half22float2(*reinterpret_cast<const __half2*>(&gb.y));
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
            *reinterpret_cast<uint4*>(dout + tile_chunk_off(row, 0, NET_OUT)) =
                make_uint4(*reinterpret_cast<const uint32_t*>(&d01), *reinterpret_cast<const uint32_t*>(&d23), 0u, 0u);
            *reinterpret_cast<uint4*>(dout + tile_chunk_off(row, 1, NET_OUT)) = make_uint4(0u, 0u, 0u, 0u);
        }
        // ---- output layer: dW_8^T [in][out] += X_8^T . dOut ;  D = dOut . W_8   (W_7 is still in the other slot)
        FB_RUN({
            if (FB_KEEP(FN_HIDDEN))
                fb_mma_wgrad<NET_OUT>(tm + FB_COL_G(FN_HIDDEN - l0), tc::smem_u32(x8), tc::smem_u32(dout), have);
            fb_mma_dgrad<NET_OUT>(tm + FB_COL_D, tc::smem_u32(dout), tc::smem_u32(wb0));
        }, (void)0);
        fb_masked_row(t_row + FB_COL_D, work, row, half, mask[FN_HIDDEN - 1]);        // dZ_7
        // ---- hidden layers 7 .. 1: dW_l += dZ_l^T . X_l ;  D = dZ_l . W_l ;  dZ_{l-1} = D masked by (X_l > 0).
        // W_{l-1} is loaded under the MMAs into the slot of W_{l+1} (done with); for l = 1 that fetches W_0 for
        // the next tile's forward
        // A launch that keeps layers l0 .. l0 + 2 needs the chain only down to dZ_{l0}: below its lowest layer
        // nothing is computed (launches 3 / 6 stop after 6 / 3 of the 9 backward phases).  Its last phase is the
        // weight gradient of layer l0 alone, under which W_0 is fetched for the next tile's forward.
#pragma unroll
        for (int l = FN_HIDDEN - 1; l >= 1; l--) {
            if (l < l0) break;
            uint8_t* wc = (l & 1) ? wb1 : wb0;
            uint8_t* wn = (l & 1) ? wb0 : wb1;
            const bool last = l == l0;
            FB_RUN({
                if (FB_KEEP(l))
                    fb_mma_wgrad<FN_W>(tm + FB_COL_G(l - l0), tc::smem_u32(work), tc::smem_u32(FB_X(l)), have);
                if (!last) fb_mma_dgrad<FN_W>(tm + FB_COL_D, tc::smem_u32(work), tc::smem_u32(wc));
            }, {
                if (last) fb_load_weights(wb0, weights, 0, tid);
                else if (l - 1 != l0 || l0 == 0) fb_load_weights(wn, weights, l - 1, tid);
            });
            if (!last) fb_masked_row(t_row + FB_COL_D, work, row, half, mask[l - 1]);  // dZ_{l-1}
        }
        // ---- layer 0: dW_0 += dZ_0^T . X_0 (nothing flows into the parameter-free encoding)
        if (FB_KEEP(0))
            FB_RUN(fb_mma_wgrad<FN_W>(tm + FB_COL_G(0 - l0), tc::smem_u32(work), tc::smem_u32(FB_X(0)), have), (void)0);
        have = true;
    }
    // ---- the three weight-gradient accumulators leave TMEM once (M = 128: row r in TMEM lane r)
    tc::fence_before_thread_sync();
    __syncthreads();
    tc::fence_after_thread_sync();
    if (have) {
#pragma unroll 1
        for (int i = 0; i < 3; i++) {
            const int l = l0 + i;
            if (l > FN_HIDDEN) break;
            float* gw = grad_weights + fn_layer_src(l);
            if (l == FN_HIDDEN) {        // transposed accumulator [in = row][out j]
                if (half == 0) {
                    uint32_t r[16];
                    tc::tmem_ld_x16(t_row + FB_COL_G(i), r);
                    tc::tmem_ld_wait();
#pragma unroll
                    for (int j = 0; j < NET_OUT; j++) atomicAdd(gw + j * FN_W + row, __uint_as_float(r[j]));
                }
            } else {
                const int K = fn_layer_k(l);
#pragma unroll 1
                for (int qq = 0; qq < FN_W / 32; qq++) {
                    const int q = 4 * half + qq;
                    uint32_t r[16];
                    tc::tmem_ld_x16(t_row + FB_COL_G(i) + q * 16, r);
                    tc::tmem_ld_wait();
#pragma unroll
                    for (int j = 0; j < 16; j++)
                        if (q * 16 + j < K) atomicAdd(gw + (size_t)row * K + q * 16 + j, __uint_as_float(r[j]));
                }
            }
        }
    }
    tc::fence_before_thread_sync();
    __syncthreads();
    tc::fence_after_thread_sync();
    if (warp == 0) tc::tmem_dealloc<512>(tm);
}

extern "C" int rtxnerf_freqnet_backward(const uint16_t* d_weights, int out_activation, const float* d_samples,
                                        int sample_stride, const uint16_t* d_dl_dradiance, const long long* d_count,
                                        long long count_scale, long long n_max, float* d_grad_weights, void* stream)
{
    if (!d_weights || !d_samples || !d_dl_dradiance || !d_grad_weights) return RTXNERF_EINVAL;
    if (n_max < 0 || sample_stride < 5 || count_scale < 1) return RTXNERF_EINVAL;
    if (out_activation != 0 && out_activation != 1) return RTXNERF_ECONFIG;
    if (((uintptr_t)d_weights & 15u) || ((uintptr_t)d_dl_dradiance & 7u)) return RTXNERF_EINVAL;
    if (n_max == 0) return RTXNERF_OK;
    static RtxPerDeviceOnce attr;
    RTX_CUDA_TRY((cudaError_t)rtx_set_max_smem(attr, freqnet_backward_kernel, FbSmem::BYTES));
    const long long tiles = (n_max + NET_TILE - 1) / NET_TILE;
    const long long blocks = tiles < RTX_NUM_SMS ? tiles : RTX_NUM_SMS;      // one CTA per SM owns all 512 TMEM columns
    for (int l0 = 0; l0 < FN_LAYERS; l0 += 3)
        freqnet_backward_kernel<<<(unsigned)blocks, FB_THREADS, FbSmem::BYTES, (cudaStream_t)stream>>>(
            d_weights, d_samples, sample_stride, out_activation, d_dl_dradiance, d_count, count_scale, n_max, l0,
            d_grad_weights);
    return rtx_launch_status(3);
}
