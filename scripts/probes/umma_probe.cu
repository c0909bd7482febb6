// umma_probe.cu — development probe (not part of the library): checks on the GPU
//  (1) MN-major shared-memory descriptors over the SAME core-matrix tile layout the forward
//      kernels use K-major (LBO/SBO swapped), i.e. D[M=feat][N=feat] = sum_s A[s][m] * B[s][n];
//  (2) which TMEM lanes an M=64 accumulator occupies.
// Build: nvcc -std=c++17 -gencode arch=compute_100a,code=sm_100a -I rtx_nerf_b200/csrc scripts/probes/umma_probe.cu -o /tmp/umma_probe
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cuda_fp16.h>
#include "tcgen05.cuh"

__device__ __forceinline__ uint32_t chunk_off(int r, int chunk, int K) { return (uint32_t)((r >> 3) * (K * 16) + chunk * 128 + (r & 7) * 16); }

__host__ __device__ constexpr uint32_t idesc(int M, int N, int a_mn, int b_mn)
{
    return (1u << 4) | ((uint32_t)a_mn << 15) | ((uint32_t)b_mn << 16) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}

// A: [128 samples][FA feats], B: [128 samples][FB feats] fp16 row-major in global
template <int M, int FA, int FB>
__global__ void probe(const __half* A, const __half* B, float* D /* [128 lanes][FB] */)
{
    extern __shared__ __align__(128) uint8_t smem[];
    uint8_t* sa = smem;
    uint8_t* sb = smem + 128 * FA * 2;
    uint64_t* bar = reinterpret_cast<uint64_t*>(sb + 128 * FB * 2);
    uint32_t* slot = reinterpret_cast<uint32_t*>(bar + 1);
    const int tid = threadIdx.x, warp = tid >> 5;
    for (int c = 0; c < FA / 8; c++)
        *reinterpret_cast<uint4*>(sa + chunk_off(tid, c, FA)) = *reinterpret_cast<const uint4*>(A + tid * FA + c * 8);
    for (int c = 0; c < FB / 8; c++)
        *reinterpret_cast<uint4*>(sb + chunk_off(tid, c, FB)) = *reinterpret_cast<const uint4*>(B + tid * FB + c * 8);
    if (tid == 0) { tc::mbar_init(bar, 1); tc::fence_mbar_init(); }
    if (warp == 0) tc::tmem_alloc<128>(slot);
    tc::fence_proxy_async_smem();
    tc::fence_before_thread_sync();
    __syncthreads();
    tc::fence_after_thread_sync();
    const uint32_t tm = *slot;
    if (tid == 0) {
        // MN-major: SBO = stride between 8-feature groups = 128 B, LBO = stride between 8-sample groups
        for (int ks = 0; ks < 128 / 16; ks++) {
            const uint64_t ad = tc::smem_desc(tc::smem_u32(sa) + ks * 2 * (FA * 16), FA * 16, 128);
            const uint64_t bd = tc::smem_desc(tc::smem_u32(sb) + ks * 2 * (FB * 16), FB * 16, 128);
            tc::mma_f16_ss(tm, ad, bd, idesc(M, FB, 1, 1), ks > 0);
        }
        tc::mma_commit(bar);
    }
    tc::mbar_wait(bar, 0);
    tc::fence_after_thread_sync();
    for (int q = 0; q < FB / 16; q++) {
        uint32_t r[16];
        tc::tmem_ld_x16(tm + ((uint32_t)(warp * 32) << 16) + q * 16, r);
        tc::tmem_ld_wait();
        for (int j = 0; j < 16; j++) D[tid * FB + q * 16 + j] = __uint_as_float(r[j]);
    }
    tc::fence_before_thread_sync();
    __syncthreads();
    if (warp == 0) tc::tmem_dealloc<128>(tm);
}

template <int M, int FA, int FB>
int run()
{
    std::vector<__half> A(128 * FA), B(128 * FB);
    std::vector<float> Af(128 * FA), Bf(128 * FB);
    srand(1);
    for (size_t i = 0; i < A.size(); i++) { float v = (rand() % 17 - 8) / 8.0f; A[i] = __float2half(v); Af[i] = v; }
    for (size_t i = 0; i < B.size(); i++) { float v = (rand() % 13 - 6) / 4.0f; B[i] = __float2half(v); Bf[i] = v; }
    __half *dA, *dB; float* dD;
    cudaMalloc(&dA, A.size() * 2); cudaMalloc(&dB, B.size() * 2); cudaMalloc(&dD, 128 * FB * 4);
    cudaMemcpy(dA, A.data(), A.size() * 2, cudaMemcpyHostToDevice);
    cudaMemcpy(dB, B.data(), B.size() * 2, cudaMemcpyHostToDevice);
    cudaMemset(dD, 0xff, 128 * FB * 4);
    const int smem = 128 * FA * 2 + 128 * FB * 2 + 64;
    cudaFuncSetAttribute(probe<M, FA, FB>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    probe<M, FA, FB><<<1, 128, smem>>>(dA, dB, dD);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("M=%d FA=%d FB=%d: CUDA error %s\n", M, FA, FB, cudaGetErrorString(e)); return 1; }
    std::vector<float> D(128 * FB);
    cudaMemcpy(D.data(), dD, D.size() * 4, cudaMemcpyDeviceToHost);
    // reference: R[m][n] = sum_s A[s][m] * B[s][n], m < min(M, FA)
    int rows = M < FA ? M : FA;
    std::vector<int> lane_of(rows, -1);
    int bad = 0;
    for (int m = 0; m < rows; m++) {
        std::vector<float> ref(FB);
        for (int n = 0; n < FB; n++) { float s = 0; for (int k = 0; k < 128; k++) s += Af[k * FA + m] * Bf[k * FB + n]; ref[n] = s; }
        for (int lane = 0; lane < 128; lane++) {
            bool ok = true;
            for (int n = 0; n < FB; n++) if (fabsf(D[lane * FB + n] - ref[n]) > 1e-2f) { ok = false; break; }
            if (ok) { lane_of[m] = lane; break; }
        }
        if (lane_of[m] < 0) bad++;
    }
    printf("M=%d FA=%d FB=%d: %d/%d result rows found; row->lane:", M, FA, FB, rows - bad, rows);
    for (int m = 0; m < rows; m += (rows > 16 ? 8 : 1)) printf(" %d->%d", m, lane_of[m]);
    printf(" ... %d->%d\n", rows - 1, lane_of[rows - 1]);
    return bad;
}

int main()
{
    int bad = 0;
    bad += run<64, 64, 64>();
    bad += run<64, 64, 32>();
    bad += run<64, 64, 16>();
    bad += run<128, 64, 64>();   // M=128 over a 64-feature A tile: rows 64..127 read past the tile (ignored)
    printf(bad ? "PROBE FAILED\n" : "PROBE OK\n");
    return bad ? 1 : 0;
}
