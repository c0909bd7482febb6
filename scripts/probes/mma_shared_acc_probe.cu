// mma_shared_acc_probe.cu — do tcgen05 MMAs issued by DIFFERENT threads of a CTA into the SAME TMEM accumulator
// lose updates?  1, 2 and 4 issuing warps each add `per` products of all-ones fp16 tiles (K = 16) into one
// accumulator that was zeroed with tcgen05.st; every element must end at issuers * per * 16 exactly (fp32 is exact
// up to 2^24).  Build: nvcc -gencode arch=compute_100a,code=sm_100a -o scripts/probes/mma_shared_acc_probe <this file>
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include "../../rtx_nerf_b200/csrc/tcgen05.cuh"

template <int M, int N>
__global__ void __launch_bounds__(160, 1) probe(int issuers, int per, float* out)
{
    extern __shared__ __align__(128) uint8_t smem[];
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + 96 * 1024);
    uint32_t* slot = reinterpret_cast<uint32_t*>(bars + 8);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int i = threadIdx.x; i < 96 * 1024 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0x3C003C00u;   // 1.0
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
    if (warp < 4) {      // zero the accumulator: 128 lanes x N columns
        uint32_t z[16];
        for (int i = 0; i < 16; i++) z[i] = 0u;
        for (int c = 0; c < N; c += 16) tc::tmem_st_x16(tm + ((uint32_t)(warp * 32) << 16) + c, z);
        tc::tmem_st_wait();
    }
    tc::fence_before_thread_sync();
    __syncthreads();
    tc::fence_after_thread_sync();
    if (warp < issuers) {
        if (tc::elect_one()) {
            constexpr uint32_t id = tc::idesc_f16_f32(M, N);
            const uint32_t a = tc::smem_u32(smem), b = tc::smem_u32(smem + 48 * 1024);
            const uint64_t da = tc::smem_desc(a, 128, 1024), db = tc::smem_desc(b, 128, 1024);
#pragma unroll 1
            for (int i = 0; i < per; i += 8) {
#pragma unroll
                for (int k = 0; k < 8; k++)
                    tc::mma_f16_ss(tm, tc::smem_desc_advance(da, k * 256), tc::smem_desc_advance(db, k * 256), id, 1u);
            }
            tc::mma_commit(bars + warp);
            tc::mbar_wait(bars + warp, 0);
        }
    }
    tc::fence_before_thread_sync();
    __syncthreads();
    tc::fence_after_thread_sync();
    if (warp < 4) {
        uint32_t r[16];
        float mn = 3.0e38f, mx = -3.0e38f;
        for (int c = 0; c < N; c += 16) {
            tc::tmem_ld_x16(tm + ((uint32_t)(warp * 32) << 16) + c, r);
            tc::tmem_ld_wait();
            for (int i = 0; i < 16; i++) { mn = fminf(mn, __uint_as_float(r[i])); mx = fmaxf(mx, __uint_as_float(r[i])); }
        }
        const bool used = M == 128 || lane < 16;      // M = 64: rows live in lanes 0..15 of each quarter
        out[2 * threadIdx.x] = used ? mn : -1.0f;
        out[2 * threadIdx.x + 1] = used ? mx : -1.0f;
    }
    tc::fence_before_thread_sync();
    __syncthreads();
    if (warp == 4) tc::tmem_dealloc<512>(tm);
}

template <int M, int N>
static void run(const char* name)
{
    float* d_out;
    cudaMalloc(&d_out, sizeof(float) * 256);
    cudaFuncSetAttribute(probe<M, N>, cudaFuncAttributeMaxDynamicSharedMemorySize, 97 * 1024);
    for (int issuers = 1; issuers <= 4; issuers *= 2) {
        const int per = 4096;
        for (int rep = 0; rep < 3; rep++) {
            probe<M, N><<<1, 160, 97 * 1024>>>(issuers, per, d_out);
            cudaError_t e = cudaDeviceSynchronize();
            if (e != cudaSuccess) { printf("%s: %s\n", name, cudaGetErrorString(e)); return; }
            float h[256];
            cudaMemcpy(h, d_out, sizeof(h), cudaMemcpyDeviceToHost);
            float mn = 3.0e38f, mx = -3.0e38f;
            for (int t = 0; t < 128; t++) {
                if (h[2 * t] < 0.0f) continue;
                mn = h[2 * t] < mn ? h[2 * t] : mn;
                mx = h[2 * t + 1] > mx ? h[2 * t + 1] : mx;
            }
            const float want = (float)issuers * per * 16.0f;
            printf("%-14s issuers %d run %d: accumulator min %.0f max %.0f, expected %.0f  %s\n", name, issuers, rep, mn, mx,
                   want, (mn == want && mx == want) ? "OK" : "UPDATES LOST");
        }
    }
    cudaFree(d_out);
}

int main()
{
    run<64, 64>("M64  N64  K16");
    run<64, 16>("M64  N16  K16");
    run<128, 64>("M128 N64  K16");
    return 0;
}
