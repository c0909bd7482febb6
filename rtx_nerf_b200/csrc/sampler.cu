// sampler.cu — per-segment sample generation for sm_100a.
//
// Replaces generate_samples / launchSampler (sampler/sampler.cu:14-131).  Same outputs and
// layouts: samples[((indices[x]+j)*S+i)*5 + {0..4}] = x,y,z,theta,phi and
// t_vals[(indices[x]+j)*S+i] (sampler.cu:60-66), with the reference's quirks kept (Q12:
// t_vals is segment-local, (i+1)/S in REGULAR mode; Q14: every ray draws the same
// minstd_rand sequence because the generator is a by-value kernel argument).
//
// The reference runs one thread per ray, serial over hits x 32 samples, with stride-5
// scalar stores (~1/8 sector efficiency).  Here one warp owns one ray: lanes walk the
// ray's samples, which are contiguous in both output arrays, the 5-float records are
// transposed through shared memory and leave as 128-bit coalesced stores.
// HBM-bound: 24 B read per segment, 24 B written per sample.
#include "common.cuh"

#define SAMPLER_WARPS 8

// thrust::minstd_rand = LCG(48271, 0, 2^31-1), default seed 1: the n-th draw is 48271^n mod m.
__device__ __forceinline__ uint32_t minstd_draw(uint32_t n)
{
    const uint64_t m = 2147483647ull;
    uint64_t r = 1, b = 48271ull;
    while (n) {
        if (n & 1u) r = (r * b) % m;
        b = (b * b) % m;
        n >>= 1;
    }
    return (uint32_t)r;
}
// thrust::uniform_real_distribution<float>(a,b) on minstd_rand (sampler.cu:69-70,83-84):
//   r = float(x - min) / (1 + float(max - min));  return r*(b-a) + a   [fma under nvcc]
__device__ __forceinline__ float uniform_real(uint32_t draw_index, float a, float b)
{
    float r = (float)(minstd_draw(draw_index) - 1u);
    r = rn_div(r, rn_add(1.0f, (float)(2147483646u - 1u)));
    return rn_fma(r, rn_sub(b, a), a);
}

// Counter-based draw in [0,1) for the seeded modes (oracle: orc_counter_uniform):
// splitmix64 of (seed, global sample index), top 24 bits.  Every sample has its own draw and a
// launch is reproducible from (seed, index) alone.
__device__ __forceinline__ float counter_uniform(unsigned long long seed, unsigned long long index)
{
    unsigned long long z = seed + (index + 1ull) * 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    z = z ^ (z >> 31);
    return (float)(z >> 40) * (1.0f / 16777216.0f);
}

// i repeated additions of inv_s starting from `from` (sampler.cu:65,97: t += 1.0f/S)
__device__ __forceinline__ float accumulate_steps(float from, float inv_s, int i, bool exact_pow2)
{
    if (exact_pow2) return rn_fma((float)i, inv_s, from);  // exact: no rounding ever occurs
    float t = from;
    for (int k = 0; k < i; k++) t = rn_add(t, inv_s);
    return t;
}

__global__ void __launch_bounds__(SAMPLER_WARPS * 32)
sampler_kernel(const float* __restrict__ start_points, const float* __restrict__ end_points,
               const float* __restrict__ view_dirs, float* __restrict__ t_vals, float* __restrict__ samples,
               int batch_size, const int* __restrict__ num_hits, const int* __restrict__ indices, int S,
               int sample_type, unsigned long long seed)
{
    __shared__ __align__(16) float tile[SAMPLER_WARPS][160];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int warps_total = gridDim.x * SAMPLER_WARPS;
    const float inv_s = 1.0f / (float)S;
    const bool pow2 = (S & (S - 1)) == 0;
    float* my_tile = tile[warp];

    for (int x = blockIdx.x * SAMPLER_WARPS + warp; x < batch_size; x += warps_total) {
        const int n_hits = num_hits[x];
        if (n_hits <= 0) continue;
        const long long first = (long long)indices[x];
        const float2 vd = reinterpret_cast<const float2*>(view_dirs)[x];
        const long long n = (long long)n_hits * S;       // samples of this ray
        const long long out0 = first * S;                // first sample index
        for (long long q0 = 0; q0 < n; q0 += 32) {
            const long long q = q0 + lane;
            const bool valid = q < n;
            float sx = 0.f, sy = 0.f, sz = 0.f, tv = 0.f;
            if (valid) {
                const int j = (int)(q / S), i = (int)(q - (long long)j * S);
                const float* a = start_points + 3 * (first + j);
                const float* b = end_points + 3 * (first + j);
                const float ax = a[0], ay = a[1], az = a[2];
                const float dx = rn_sub(b[0], ax), dy = rn_sub(b[1], ay), dz = rn_sub(b[2], az);  // :45-47
                float t;
                if (sample_type == 0) {                  // REGULAR :54-66
                    t = accumulate_steps(0.0f, inv_s, i, pow2);
                    tv = rn_add(t, inv_s);
                } else if (sample_type == 2) {           // UNIFORM :68-81 (t_vals stays t_initial = 0)
                    t = seed ? counter_uniform(seed, (unsigned long long)(out0 + q))
                             : uniform_real((uint32_t)(q + 1), 0.0f, 1.0f);
                    tv = 0.0f;
                } else {                                 // STRATIFIED_JITTERING :82-99
                    float t0 = accumulate_steps(0.0f, inv_s, i, pow2);   // t_initial after i steps
                    float t1 = accumulate_steps(inv_s, inv_s, i, pow2);  // t_final after i steps
                    // t_initial = t_final (i>0) : after the first step t_initial carries t_final's value
                    if (i > 0) t0 = accumulate_steps(inv_s, inv_s, i - 1, pow2);
                    t = seed ? rn_fma(counter_uniform(seed, (unsigned long long)(out0 + q)), rn_sub(t1, t0), t0)
                             : uniform_real((uint32_t)(q + 1), t0, t1);
                    tv = t0;
                }
                sx = rn_fma(t, dx, ax);                  // :57-59 (fma under nvcc -fmad)
                sy = rn_fma(t, dy, ay);
                sz = rn_fma(t, dz, az);
            }
            const long long o = out0 + q0;               // first sample of this 32-chunk
            if (valid) t_vals[o + lane] = tv;            // coalesced 128 B
            // 5-float records: transpose through shared memory (stride 5 is conflict-free)
            my_tile[lane * 5 + 0] = sx;
            my_tile[lane * 5 + 1] = sy;
            my_tile[lane * 5 + 2] = sz;
            my_tile[lane * 5 + 3] = vd.x;
            my_tile[lane * 5 + 4] = vd.y;
            __syncwarp();
            const int cnt = (int)min((long long)32, n - q0);   // samples in this chunk
            float* dst = samples + 5 * o;
            if ((o & 3) == 0) {                          // 20*o bytes is 16-byte aligned
                const int nvec = (cnt * 5) >> 2;
                for (int v = lane; v < nvec; v += 32)
                    st_stream_f4(reinterpret_cast<float4*>(dst) + v, reinterpret_cast<const float4*>(my_tile)[v]);
                for (int e = (nvec << 2) + lane; e < cnt * 5; e += 32) dst[e] = my_tile[e];
            } else {
                for (int e = lane; e < cnt * 5; e += 32) dst[e] = my_tile[e];
            }
            __syncwarp();
        }
    }
}

extern "C" int rtxnerf_launch_sampler(const float* d_start_points, const float* d_end_points,
                                      const float* d_view_dirs, float* d_t_vals, float* d_sampled_points,
                                      int batch_size, int grid_res, const int* d_num_hits, const int* d_indices,
                                      int sample_type, int samples_per_segment, void* stream)
{
    return rtxnerf_launch_sampler_seeded(d_start_points, d_end_points, d_view_dirs, d_t_vals, d_sampled_points,
                                         batch_size, grid_res, d_num_hits, d_indices, sample_type,
                                         samples_per_segment, 0ull, stream);
}

extern "C" int rtxnerf_launch_sampler_seeded(const float* d_start_points, const float* d_end_points,
                                             const float* d_view_dirs, float* d_t_vals, float* d_sampled_points,
                                             int batch_size, int grid_res, const int* d_num_hits,
                                             const int* d_indices, int sample_type, int samples_per_segment,
                                             unsigned long long seed, void* stream)
{
    (void)grid_res;  // unused by the reference too (sampler.cu:19)
    if (!d_start_points || !d_end_points || !d_view_dirs || !d_t_vals || !d_sampled_points || !d_num_hits ||
        !d_indices)
        return RTXNERF_EINVAL;
    if (batch_size < 0 || samples_per_segment < 1 || sample_type < 0 || sample_type > 2) return RTXNERF_EINVAL;
    // the kernel moves whole 16-byte vectors (sample rows) and 8-byte view directions: a sub-buffer that breaks
    // that alignment is refused here instead of faulting on the device
    if (((uintptr_t)d_sampled_points & 15u) || ((uintptr_t)d_view_dirs & 7u)) return RTXNERF_EINVAL;
    if (batch_size == 0) return RTXNERF_OK;
    long long blocks = ((long long)batch_size + SAMPLER_WARPS - 1) / SAMPLER_WARPS;
    const long long max_blocks = (long long)RTX_NUM_SMS * 64;
    if (blocks > max_blocks) blocks = max_blocks;
    sampler_kernel<<<(unsigned)blocks, SAMPLER_WARPS * 32, 0, (cudaStream_t)stream>>>(
        d_start_points, d_end_points, d_view_dirs, d_t_vals, d_sampled_points, batch_size, d_num_hits, d_indices,
        samples_per_segment, sample_type, seed);
    return rtx_launch_status();
}
