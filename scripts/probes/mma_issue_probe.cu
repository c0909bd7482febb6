// mma_issue_probe.cu — how fast can ONE thread issue small tcgen05 MMAs, and does a second / fourth issuing
// thread of the same CTA add to the rate?  (Design question behind mlp_backward: 43 small MMAs per 128-sample tile.)
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -o scripts/probes/mma_issue_probe scripts/probes/mma_issue_probe.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include "../../rtx_nerf_b200/csrc/tcgen05.cuh"

template <int M, int N, int AMN, int BMN>
__global__ void __launch_bounds__(160, 1) probe(int issuers, int per_issuer, long long* out)
{
    extern __shared__ __align__(128) uint8_t smem[];
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + 96 * 1024);
    uint32_t* slot = reinterpret_cast<uint32_t*>(bars + 8);
    const int warp = threadIdx.x >> 5;
    for (int i = threadIdx.x; i < 96 * 1024 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0u;
    if (threadIdx.x == 0) {
        for (int i = 0; i < 4; i++) tc::mbar_init(bars + i, 1);
        tc::fence_mbar_init();
    }
    if (warp == 4) tc::tmem_alloc<512>(slot);
    tc::fence_proxy_async_smem();
    tc::fence_before_thread_sync();
    __syncthreads();
    tc::fence_after_thread_sync();
    const uint32_t tm = *slot;
    long long t_issue = 0, t_done = 0;
    if (warp < issuers) {
        if (tc::elect_one()) {
            constexpr uint32_t id = tc::idesc_f16_f32_major(M, N, AMN, BMN);
            const uint32_t a = tc::smem_u32(smem), b = tc::smem_u32(smem + 48 * 1024);
            // K-major: [rows][64] tile, LBO 128, SBO 1024, next K step +256 B.  MN-major: the [128 samples][rows]
            // tile of the backward kernels (LBO = rows * 16, SBO 128, next K step + 2 * rows * 16)
            const uint64_t da = AMN ? tc::smem_desc(a, M * 16, 128) : tc::smem_desc(a, 128, 1024);
            const uint64_t db = BMN ? tc::smem_desc(b, N * 16, 128) : tc::smem_desc(b, 128, 1024);
            const uint32_t sa = AMN ? 2 * M * 16 : 256, sb = BMN ? 2 * N * 16 : 256;
            const uint32_t d = tm + 128 * warp;      // each issuer its own accumulator columns
            const long long t0 = clock64();
#pragma unroll 1
            for (int i = 0; i < per_issuer; i += 8) {
#pragma unroll
                for (int k = 0; k < 8; k++)
                    tc::mma_f16_ss(d, tc::smem_desc_advance(da, k * sa), tc::smem_desc_advance(db, k * sb), id, 1u);
            }
            tc::mma_commit(bars + warp);
            const long long t1 = clock64();
            tc::mbar_wait(bars + warp, 0);
            const long long t2 = clock64();
            t_issue = t1 - t0;
            t_done = t2 - t0;
            out[2 * (blockIdx.x * 4 + warp)] = t_issue;
            out[2 * (blockIdx.x * 4 + warp) + 1] = t_done;
        }
    }
    tc::fence_before_thread_sync();
    __syncthreads();
    tc::fence_after_thread_sync();
    if (warp == 4) tc::tmem_dealloc<512>(tm);
}

template <int M, int N, int AMN = 0, int BMN = 0>
static void run(const char* name, int blocks)
{
    long long* d_out;
    cudaMalloc(&d_out, sizeof(long long) * 8 * blocks);
    cudaFuncSetAttribute(probe<M, N, AMN, BMN>, cudaFuncAttributeMaxDynamicSharedMemorySize, 97 * 1024);
    for (int issuers = 1; issuers <= 4; issuers *= 2) {
        const int per = 4096;
        cudaMemset(d_out, 0, sizeof(long long) * 8 * blocks);
        probe<M, N, AMN, BMN><<<blocks, 160, 97 * 1024>>>(issuers, per, d_out);
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) { printf("%s: %s\n", name, cudaGetErrorString(e)); return; }
        long long h[8];
        cudaMemcpy(h, d_out, sizeof(h), cudaMemcpyDeviceToHost);
        long long worst_issue = 0, worst_done = 0;
        for (int w = 0; w < issuers; w++) { worst_issue = h[2 * w] > worst_issue ? h[2 * w] : worst_issue; worst_done = h[2 * w + 1] > worst_done ? h[2 * w + 1] : worst_done; }
        printf("%-22s issuers %d: issue %.1f cycles/MMA/thread, all done after %.1f cycles per MMA (%d MMAs in total; tensor floor %d cycles each)\n",
               name, issuers, (double)worst_issue / per, (double)worst_done / (per * issuers), per * issuers,
               (M > 128 ? M : 128) * N / 256);
    }
    cudaFree(d_out);
}

int main()
{
    run<64, 16>("M64  N16  K16", 1);
    run<64, 64>("M64  N64  K16", 1);
    run<128, 16>("M128 N16  K16", 1);
    run<128, 32>("M128 N32  K16", 1);
    run<128, 64>("M128 N64  K16", 1);
    run<128, 128>("M128 N128 K16", 1);
    run<128, 64, 0, 1>("dgrad M128 N64 B-MN", 1);
    run<128, 32, 0, 1>("dgrad M128 N32 B-MN", 1);
    run<64, 64, 1, 1>("wgrad M64 N64 MN/MN", 1);
    run<64, 32, 1, 1>("wgrad M64 N32 MN/MN", 1);
    run<64, 16, 1, 1>("wgrad M64 N16 MN/MN", 1);
    run<128, 64, 1, 1>("wgrad M128 N64 MN/MN", 1);
    run<128, 128, 1, 1>("wgrad M128 N128 MN/MN", 1);
    return 0;
}
