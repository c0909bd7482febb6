// grid_update.cu — occupancy-grid rebuild from densities on the device, and a headless image
// writer.  SURVEY §8 rows f3 / f4.
//
// The reference only has the static all-cells list of make_grid (main.cu:154-174, with its own
// "TODO: make this a CUDA kernel", :153) and rebuilds nothing during training; its OptiX GAS
// would have to be rebuilt on the host.  Here the grid is a bitfield, so an update is one pass:
//   rtxnerf_grid_cell_samples        [G^3][5] rows at the cell centres (the network's input rows)
//   rtxnerf_grid_update_from_density ema = max(ema*decay, sigma); bit = ema > threshold  (the
//                                    instant-ngp update rule; decay = 0 / ema = NULL: plain threshold)
// One thread per cell, one warp per 32-bit word (bit p&31 of word p>>5 = lane), ballot packs.
// stb_image_write is vendored but unused by the reference (main.cu:19-21); rtxnerf_write_ppm is the
// headless stand-in: fp32 RGB in [0,1] -> binary P6.
#include <stdio.h>

#include <vector>

#include "common.cuh"

__global__ void grid_cell_samples_kernel(int G, float* __restrict__ rows)
{
    const long long p = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const long long n = (long long)G * G * G;
    if (p >= n) return;
    const int z = (int)(p % G), y = (int)((p / G) % G), x = (int)(p / ((long long)G * G));
    const float L = 2.0f / (float)G, h = 1.0f / (float)G;   // cell size and half of it (exact: G is a power of two)
    float* r = rows + 5 * p;
    r[0] = rn_add(rn_fma((float)x, L, -1.0f), h);            // main.cu:157-167: min = -1 + i*(2/G)
    r[1] = rn_add(rn_fma((float)y, L, -1.0f), h);
    r[2] = rn_add(rn_fma((float)z, L, -1.0f), h);
    r[3] = 0.0f;
    r[4] = 0.0f;
}

__global__ void grid_update_kernel(long long n_cells, const float* __restrict__ sigma, int stride, float decay,
                                   float* __restrict__ ema, float threshold, uint32_t* __restrict__ bits)
{
    const long long p = (long long)blockIdx.x * blockDim.x + threadIdx.x;   // blockDim is a multiple of 32
    float v = 0.0f;
    bool occ = false;
    if (p < n_cells) {
        v = sigma[p * stride];
        if (ema) {
            v = fmaxf(rn_mul(ema[p], decay), v);
            ema[p] = v;
        }
        occ = v > threshold;
    }
    const unsigned word = __ballot_sync(0xffffffffu, occ);
    if ((threadIdx.x & 31) == 0 && p < n_cells) bits[p >> 5] = word;
}

extern "C" int rtxnerf_grid_cell_samples(int G, float* d_rows5, void* stream)
{
    if (!d_rows5) return RTXNERF_EINVAL;
    if (G < 2 || G > 1024 || (G & (G - 1))) return RTXNERF_EGRID;
    const long long n = (long long)G * G * G;
    grid_cell_samples_kernel<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(G, d_rows5);
    return rtx_launch_status();
}

extern "C" int rtxnerf_grid_update_from_density(rtxnerf_grid* g, const float* d_sigma, int sigma_stride, float decay,
                                                float* d_ema, float threshold, void* stream)
{
    if (!g || !d_sigma || sigma_stride < 1) return RTXNERF_EINVAL;
    if (!(decay >= 0.0f && decay <= 1.0f)) return RTXNERF_EINVAL;
    const long long n = (long long)g->G * g->G * g->G;
    grid_update_kernel<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(n, d_sigma, sigma_stride, decay,
                                                                                    d_ema, threshold, g->d_bits);
    const int rc = rtx_launch_status();
    if (rc) return rc;
    return rtx_grid_rebuild_bricks(g, (cudaStream_t)stream);
}

extern "C" int rtxnerf_grid_read_bits(const rtxnerf_grid* g, uint32_t* bits_host)
{
    if (!g || !bits_host) return RTXNERF_EINVAL;
    RTX_CUDA_TRY(cudaMemcpy(bits_host, g->d_bits, g->n_words * sizeof(uint32_t), cudaMemcpyDeviceToHost));
    return RTXNERF_OK;
}

extern "C" int rtxnerf_write_ppm(const char* path, const float* pixels_host, int width, int height)
{
    if (!path || !pixels_host || width < 1 || height < 1) return RTXNERF_EINVAL;
    FILE* f = fopen(path, "wb");
    if (!f) return RTXNERF_EINVAL;
    fprintf(f, "P6\n%d %d\n255\n", width, height);
    std::vector<unsigned char> row((size_t)width * 3);
    for (int y = 0; y < height; y++) {
        const float* p = pixels_host + (size_t)y * width * 3;   // ray_idx = x + y*W (optixPrograms.cu:45)
        for (int i = 0; i < width * 3; i++) {
            float v = p[i];
            v = v != v ? 0.0f : (v < 0.0f ? 0.0f : (v > 1.0f ? 1.0f : v));
            row[(size_t)i] = (unsigned char)(v * 255.0f + 0.5f);
        }
        if (fwrite(row.data(), 1, row.size(), f) != row.size()) {
            fclose(f);
            return RTXNERF_EINVAL;
        }
    }
    fclose(f);
    return RTXNERF_OK;
}
