// batch.cu — seeded ray shuffle + batch assembly on the device.
//
// Replaces the reference's training-loop front end (main.cu:612-676): std::random_shuffle of
// the host ray payloads, the per-iteration host loops that collect num_hits / view_dir / gt
// pixel, three H2D copies, thrust::reduce + exclusive_scan, the host gather of start/end points
// into freshly malloc'ed arrays and two more H2D copies.  Here the march results of ALL views
// stay resident in HBM (a "ray store": 100 views x 800x800 rays x ~14 hits x 24 B = 22 GB of
// 180 GB) and a batch is assembled by three small kernels:
//   1. batch_count: batch slot k -> ray id through a stateless seeded bijection (no permutation
//      array, no sort), per-ray fields gathered, one partial hit sum per CTA;
//   2. batch_scan: exclusive scan of the CTA sums (one CTA);
//   3. batch_gather: CSR index = CTA offset + block-local scan; the CTA's hit records are
//      contiguous in the batch arrays, so threads take batch slots, find the owning ray by a
//      binary search over the CTA's scan in shared memory and copy the records with
//      slot-consecutive (coalesced) stores.
// HBM-bound: 36 B per ray + 2 x (24..32) B per hit.
#include "common.cuh"

#define BATCH_BLOCK 256

// ------------------------------------------------------------------- the seeded bijection
// perm(i), 0 <= i < n: 4-round Feistel network over the smallest even bit width 2h with
// 2^(2h) >= n, cycle-walked into [0, n) (at most 4x oversampling, so < 4 walks expected).
// Round function: murmur3 finaliser of (right half ^ round key); round keys = splitmix64
// stream of the seed.  std::random_shuffle's generator is implementation-defined, so the
// contract here is "a seeded permutation of the rays", restated in oracle/oracle.py
// (ray_permutation) bit for bit.
struct PermKey {
    unsigned int k[4];
    unsigned int half_bits;
    unsigned long long n;
};

static inline unsigned long long splitmix64_host(unsigned long long& s)
{
    unsigned long long z = (s += 0x9E3779B97F4A7C15ull);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

static PermKey make_perm_key(unsigned long long n, unsigned long long seed)
{
    PermKey key;
    unsigned int bits = 2;
    while (bits < 64 && (1ull << bits) < n) bits++;
    if (bits & 1) bits++;
    key.half_bits = bits / 2;
    key.n = n;
    unsigned long long s = seed;
    for (int r = 0; r < 4; r++) key.k[r] = (unsigned int)(splitmix64_host(s) >> 32);
    return key;
}

__device__ __forceinline__ unsigned int mix32(unsigned int h)
{
    h ^= h >> 16;
    h *= 0x85EBCA6Bu;
    h ^= h >> 13;
    h *= 0xC2B2AE35u;
    h ^= h >> 16;
    return h;
}

__device__ __forceinline__ unsigned long long perm_index(unsigned long long i, const PermKey& key)
{
    const unsigned int hb = key.half_bits;
    const unsigned int mask = hb >= 32 ? 0xFFFFFFFFu : ((1u << hb) - 1u);
    unsigned long long x = i;
    do {
        unsigned int L = (unsigned int)(x >> hb) & mask, R = (unsigned int)x & mask;
#pragma unroll
        for (int r = 0; r < 4; r++) {
            unsigned int F = mix32(R ^ key.k[r]) & mask;
            unsigned int t = L ^ F;
            L = R;
            R = t;
        }
        x = ((unsigned long long)L << hb) | R;
    } while (x >= key.n);
    return x;
}

// batch slot -> ray id: the permutation acts on groups of `group` consecutive rays (group = 1:
// the reference's per-ray shuffle; group = 32 keeps a warp's rays on adjacent pixels, which the
// ray-coherent network kernels use)
__device__ __forceinline__ long long slot_ray(long long pos, int group, const PermKey& key)
{
    unsigned long long g = (unsigned long long)pos / (unsigned)group;
    unsigned int in = (unsigned int)((unsigned long long)pos % (unsigned)group);
    return (long long)(perm_index(g, key) * (unsigned)group + in);
}

__global__ void __launch_bounds__(BATCH_BLOCK)
ray_permutation_kernel(PermKey key, int group, long long first, long long count, long long* __restrict__ out)
{
    long long k = (long long)blockIdx.x * BATCH_BLOCK + threadIdx.x;
    if (k < count) out[k] = slot_ray(first + k, group, key);
}

// ------------------------------------------------------------------- kernels
__global__ void __launch_bounds__(BATCH_BLOCK)
batch_count_kernel(rtxnerf_ray_store st, PermKey key, int group, long long first, long long n_batch,
                   int* __restrict__ num_hits, float* __restrict__ view_dirs, float* __restrict__ gt_pixels,
                   long long* __restrict__ ray_ids, long long* __restrict__ src_base, int* __restrict__ block_sums)
{
    __shared__ int warp_sums[BATCH_BLOCK / 32];
    long long k = (long long)blockIdx.x * BATCH_BLOCK + threadIdx.x;
    int nh = 0;
    if (k < n_batch) {
        long long r = slot_ray(first + k, group, key);
        nh = st.num_hits[r];
        num_hits[k] = nh;
        src_base[k] = st.view_base[r / st.rays_per_view] + st.indices[r];
        reinterpret_cast<float2*>(view_dirs)[k] = reinterpret_cast<const float2*>(st.view_dirs)[r];
        if (gt_pixels) {
            const float* p = st.pixels_gt + 3 * r;
            float* q = gt_pixels + 3 * k;
            q[0] = p[0]; q[1] = p[1]; q[2] = p[2];
        }
        if (ray_ids) ray_ids[k] = r;
    }
    int s = nh;
#pragma unroll
    for (int o = 16; o; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if ((threadIdx.x & 31) == 0) warp_sums[threadIdx.x >> 5] = s;
    __syncthreads();
    if (threadIdx.x == 0) {
        int t = 0;
#pragma unroll
        for (int w = 0; w < BATCH_BLOCK / 32; w++) t += warp_sums[w];
        block_sums[blockIdx.x] = t;
    }
}

// exclusive scan of the CTA sums, one CTA (nb = n_batch / 256: 1024 at 2^18 rays)
__global__ void __launch_bounds__(1024)
batch_scan_kernel(const int* __restrict__ block_sums, int nb, long long* __restrict__ block_offsets,
                  long long* __restrict__ total)
{
    __shared__ long long warp_tot[32];
    __shared__ long long carry_s;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) carry_s = 0;
    __syncthreads();
    for (int base = 0; base < nb; base += 1024) {
        int i = base + threadIdx.x;
        long long v = i < nb ? block_sums[i] : 0;
        long long inc = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            long long n = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += n;
        }
        if (lane == 31) warp_tot[warp] = inc;
        __syncthreads();
        if (warp == 0) {
            long long w = warp_tot[lane], winc = w;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                long long n = __shfl_up_sync(0xffffffffu, winc, o);
                if (lane >= o) winc += n;
            }
            warp_tot[lane] = winc - w;
        }
        __syncthreads();
        long long carry = carry_s;
        if (i < nb) block_offsets[i] = carry + warp_tot[warp] + inc - v;
        __syncthreads();
        if (threadIdx.x == 1023) carry_s = carry + warp_tot[31] + inc;
        __syncthreads();
    }
    if (threadIdx.x == 0) *total = carry_s;
}

__global__ void __launch_bounds__(BATCH_BLOCK)
batch_gather_kernel(rtxnerf_ray_store st, long long n_batch, const int* __restrict__ num_hits,
                    const long long* __restrict__ src_base, const long long* __restrict__ block_offsets,
                    int* __restrict__ indices, float* __restrict__ t_start, float* __restrict__ t_end,
                    float* __restrict__ start_points, float* __restrict__ end_points, float* __restrict__ seg8,
                    long long capacity)
{
    __shared__ int excl[BATCH_BLOCK + 1];      // block-local exclusive scan of the hit counts
    __shared__ long long sbase[BATCH_BLOCK];
    __shared__ int warp_sums[BATCH_BLOCK / 32];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    long long k = (long long)blockIdx.x * BATCH_BLOCK + tid;
    int nh = k < n_batch ? num_hits[k] : 0;
    int inc = warp_incl_scan_i32(nh, lane);
    if (lane == 31) warp_sums[warp] = inc;
    __syncthreads();
    int woff = 0;
#pragma unroll
    for (int w = 0; w < BATCH_BLOCK / 32; w++) woff += (w < warp) ? warp_sums[w] : 0;
    const long long boff = block_offsets[blockIdx.x];
    excl[tid] = woff + inc - nh;
    sbase[tid] = k < n_batch ? src_base[k] : 0;
    if (tid == BATCH_BLOCK - 1) excl[BATCH_BLOCK] = woff + inc;
    if (k < n_batch) indices[k] = (int)(boff + woff + inc - nh);
    __syncthreads();
    const int total = excl[BATCH_BLOCK];
    for (int s = tid; s < total; s += BATCH_BLOCK) {
        // owner = last ray with excl[ray] <= s (rays without hits share their successor's offset)
        int lo = 0, hi = BATCH_BLOCK - 1;
#pragma unroll
        for (int it = 0; it < 8; it++) {
            int mid = (lo + hi + 1) >> 1;
            if (excl[mid] <= s) lo = mid; else hi = mid - 1;
        }
        const long long src = sbase[lo] + (s - excl[lo]);
        const long long dst = boff + s;
        if (dst >= capacity) continue;
        const float* sp = st.start_points + 3 * src;
        const float* ep = st.end_points + 3 * src;
        float a0 = sp[0], a1 = sp[1], a2 = sp[2], b0 = ep[0], b1 = ep[1], b2 = ep[2];
        if (start_points) {
            float* dp = start_points + 3 * dst;
            float* dq = end_points + 3 * dst;
            dp[0] = a0; dp[1] = a1; dp[2] = a2;
            dq[0] = b0; dq[1] = b1; dq[2] = b2;
        }
        if (seg8) {      // packed 32-byte record (march.cu: store_seg8)
            float4* r = reinterpret_cast<float4*>(seg8) + 2 * dst;
            r[0] = make_float4(a0, a1, a2, b0);
            r[1] = make_float4(b1, b2, 0.0f, 0.0f);
        }
        if (t_start && st.t_start) t_start[dst] = st.t_start[src];
        if (t_end && st.t_end) t_end[dst] = st.t_end[src];
    }
}

// ------------------------------------------------------------------- C ABI
static inline size_t align256(size_t v) { return (v + 255) & ~(size_t)255; }

extern "C" size_t rtxnerf_batch_workspace_bytes(long long batch_rays)
{
    if (batch_rays < 0) return 0;
    size_t nb = (size_t)((batch_rays + BATCH_BLOCK - 1) / BATCH_BLOCK);
    return align256((size_t)batch_rays * sizeof(long long)) + align256(nb * sizeof(int)) +
           align256(nb * sizeof(long long));
}

extern "C" int rtxnerf_ray_permutation(long long n, unsigned long long seed, int group, long long first,
                                       long long count, long long* d_out, void* stream)
{
    if (n <= 0 || group <= 0 || n % group != 0 || first < 0 || count < 0 || first + count > n || (!d_out && count))
        return RTXNERF_EINVAL;
    if (count == 0) return RTXNERF_OK;
    PermKey key = make_perm_key((unsigned long long)(n / group), seed);
    ray_permutation_kernel<<<(unsigned)((count + BATCH_BLOCK - 1) / BATCH_BLOCK), BATCH_BLOCK, 0,
                             (cudaStream_t)stream>>>(key, group, first, count, d_out);
    return rtx_launch_status();
}

static int batch_assemble_impl(const rtxnerf_ray_store* store, unsigned long long seed, int group, long long first,
                               long long batch_rays, const rtxnerf_hits* batch, float* seg8, float* d_gt_pixels,
                               long long* d_ray_ids, void* workspace, size_t workspace_bytes, void* stream);

extern "C" int rtxnerf_batch_assemble(const rtxnerf_ray_store* store, unsigned long long seed, int group,
                                      long long first, long long batch_rays, const rtxnerf_hits* batch,
                                      float* d_gt_pixels, long long* d_ray_ids, void* workspace,
                                      size_t workspace_bytes, void* stream)
{
    return batch_assemble_impl(store, seed, group, first, batch_rays, batch, nullptr, d_gt_pixels, d_ray_ids, workspace,
                               workspace_bytes, stream);
}

// rtxnerf_batch_assemble that also (or, with batch->start_points == batch->end_points == NULL, only) writes the
// batch's hits as the packed 32-byte records of rtxnerf_ray_march_csr_seg.
extern "C" int rtxnerf_batch_assemble_seg(const rtxnerf_ray_store* store, unsigned long long seed, int group,
                                          long long first, long long batch_rays, const rtxnerf_hits* batch,
                                          float* d_segments8, float* d_gt_pixels, long long* d_ray_ids,
                                          void* workspace, size_t workspace_bytes, void* stream)
{
    if (!d_segments8 || ((uintptr_t)d_segments8 & 31u)) return RTXNERF_EINVAL;
    return batch_assemble_impl(store, seed, group, first, batch_rays, batch, d_segments8, d_gt_pixels, d_ray_ids,
                               workspace, workspace_bytes, stream);
}

static int batch_assemble_impl(const rtxnerf_ray_store* store, unsigned long long seed, int group, long long first,
                               long long batch_rays, const rtxnerf_hits* batch, float* seg8, float* d_gt_pixels,
                               long long* d_ray_ids, void* workspace, size_t workspace_bytes, void* stream)
{
    if (!store || !batch || !store->num_hits || !store->indices || !store->view_base || !store->view_dirs ||
        !store->start_points || !store->end_points || !batch->num_hits || !batch->indices || !batch->view_dirs ||
        !batch->total_hits)
        return RTXNERF_EINVAL;
    if ((batch->start_points == nullptr) != (batch->end_points == nullptr)) return RTXNERF_EINVAL;
    if (!batch->start_points && !seg8) return RTXNERF_EINVAL;
    if (store->n_rays <= 0 || store->rays_per_view <= 0 || store->n_rays % store->rays_per_view != 0 ||
        group <= 0 || store->n_rays % group != 0 || first < 0 || batch_rays < 0 ||
        first + batch_rays > store->n_rays || batch_rays > 0x7fffffffLL * (long long)BATCH_BLOCK)
        return RTXNERF_EINVAL;                   // whole batches only (the reference slices past the end: Q6)
    if (d_gt_pixels && !store->pixels_gt) return RTXNERF_EINVAL;
    cudaStream_t s = (cudaStream_t)stream;
    if (batch_rays == 0) {
        RTX_CUDA_TRY(cudaMemsetAsync(batch->total_hits, 0, sizeof(long long), s));
        return RTXNERF_OK;
    }
    if (!workspace || workspace_bytes < rtxnerf_batch_workspace_bytes(batch_rays)) return RTXNERF_ECAPACITY;
    const long long nb = (batch_rays + BATCH_BLOCK - 1) / BATCH_BLOCK;
    char* w = (char*)workspace;
    long long* src_base = (long long*)w;
    w += align256((size_t)batch_rays * sizeof(long long));
    int* block_sums = (int*)w;
    w += align256((size_t)nb * sizeof(int));
    long long* block_offsets = (long long*)w;
    PermKey key = make_perm_key((unsigned long long)(store->n_rays / group), seed);
    batch_count_kernel<<<(unsigned)nb, BATCH_BLOCK, 0, s>>>(*store, key, group, first, batch_rays, batch->num_hits,
                                                            batch->view_dirs, d_gt_pixels, d_ray_ids, src_base,
                                                            block_sums);
    batch_scan_kernel<<<1, 1024, 0, s>>>(block_sums, (int)nb, block_offsets, batch->total_hits);
    batch_gather_kernel<<<(unsigned)nb, BATCH_BLOCK, 0, s>>>(*store, batch_rays, batch->num_hits, src_base,
                                                             block_offsets, batch->indices, batch->t_start,
                                                             batch->t_end, batch->start_points, batch->end_points, seg8,
                                                             batch->capacity);
    return rtx_launch_status(3);
}
