// compat_driver.cu — a C++ host program written against the REFERENCE's headers and call
// sequence (main.cu:381-399 grid/GAS, :481-508 Params + optixLaunch, :631-676 CSR compaction,
// :704-712 launchSampler, :737-746 launch_volrender_cuda, :767-777 backward), compiled with
// -Icompat so that every call lands in librtxnerf_b200.so.  It dumps its buffers to a file; the
// Python test compares them with the oracle.  Usage: compat_driver G W H pose16.bin out.bin
#include <cstdio>
#include <cstdlib>
#include <vector>

#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <optix.h>
#include <optix_stubs.h>

#include "rtx/include/rtxFunctions.h"
#include "sampler/sampler.h"
#include "vol_render/vol_render.h"

// main.cu:154-174
static std::vector<OptixAabb> make_grid(int resolution)
{
    std::vector<OptixAabb> grid;
    float box = 2.0f / resolution;
    for (int x = 0; x < resolution; x++)
        for (int y = 0; y < resolution; y++)
            for (int z = 0; z < resolution; z++) {
                OptixAabb a;
                a.minX = -1.0f + x * box; a.minY = -1.0f + y * box; a.minZ = -1.0f + z * box;
                a.maxX = a.minX + box; a.maxY = a.minY + box; a.maxZ = a.minZ + box;
                grid.push_back(a);
            }
    return grid;
}

template <class T> static T* dmalloc(size_t n) { T* p = nullptr; CUDA_CHECK(cudaMalloc(&p, n * sizeof(T))); return p; }
template <class T> static void dump(FILE* f, const std::vector<T>& v)
{
    unsigned long long n = v.size();
    fwrite(&n, sizeof(n), 1, f);
    fwrite(v.data(), sizeof(T), v.size(), f);
}

int main(int argc, char** argv)
{
    if (argc < 6) return 2;
    const int G = atoi(argv[1]);
    const unsigned W = atoi(argv[2]), H = atoi(argv[3]);
    float pose[16];
    FILE* pf = fopen(argv[4], "rb");
    if (!pf || fread(pose, 4, 16, pf) != 16) return 3;
    fclose(pf);
    const float focal = 2.7778f;   // any scalar: it is an input

    cudaStream_t inference_stream;
    CUDA_CHECK(cudaStreamCreate(&inference_stream));
    RTXDataHolder* rtx_dataholder = new RTXDataHolder();
    rtx_dataholder->initContext();
    rtx_dataholder->createModule("optixPrograms.ptx");
    rtx_dataholder->createProgramGroups();
    rtx_dataholder->linkPipeline(false);
    rtx_dataholder->buildSBT();
    std::vector<OptixAabb> grid = make_grid(G);
    OptixAabb* d_aabb = rtx_dataholder->initAccelerationStructure(grid);

    const size_t R = (size_t)W * H;
    const int arr = 3 * G;                                   // main.cu:486
    float* d_look_at = dmalloc<float>(16);
    float3* d_start = dmalloc<float3>(R * arr);
    float3* d_end = dmalloc<float3>(R * arr);
    float3* d_orig = dmalloc<float3>(R);
    float* d_ts = dmalloc<float>(R * arr);
    float* d_te = dmalloc<float>(R * arr);
    int* d_nh = dmalloc<int>(R);
    float2* d_vd = dmalloc<float2>(R);
    CUDA_CHECK(cudaMemcpy(d_look_at, pose, 64, cudaMemcpyHostToDevice));
    CUDA_CHECK(cudaMemset(d_nh, 0, R * sizeof(int)));

    Params params;                                            // main.cu:481-501
    params.handle = rtx_dataholder->gas_handle;
    params.look_at = d_look_at;
    params.num_primitives = (int)grid.size();
    params.intersection_arr_size = arr;
    params.aabb = d_aabb;
    params.start_points = d_start; params.end_points = d_end; params.ray_origins = d_orig;
    params.t_start = d_ts; params.t_end = d_te; params.num_hits = d_nh; params.viewing_direction = d_vd;
    params.focal_length = focal; params.aspect_ratio = (float)W / (float)H;
    params.delta = make_float3(2.0f / G, 2.0f / G, 2.0f / G);
    params.min_point = make_float3(-1, -1, -1); params.max_point = make_float3(1, 1, 1);
    params.width = W; params.height = H;
    Params* d_param = dmalloc<Params>(1);
    CUDA_CHECK(cudaMemcpy(d_param, &params, sizeof(Params), cudaMemcpyHostToDevice));   // main.cu:503
    const OptixShaderBindingTable& sbt_ray_march = rtx_dataholder->sbt_ray_march;
    OPTIX_CHECK(optixLaunch(rtx_dataholder->pipeline_ray_march, inference_stream, reinterpret_cast<CUdeviceptr>(d_param),
                            sizeof(Params), &sbt_ray_march, W, H, 1));                  // main.cu:506-508
    CUDA_CHECK(cudaStreamSynchronize(inference_stream));

    // D2H + host compaction to CSR, as the driver does (main.cu:517-543, 631-676)
    std::vector<int> nh(R), idx(R);
    std::vector<float3> hs(R * arr), he(R * arr);
    std::vector<float2> vd(R);
    std::vector<float> te(R * arr);
    CUDA_CHECK(cudaMemcpy(nh.data(), d_nh, R * 4, cudaMemcpyDeviceToHost));
    CUDA_CHECK(cudaMemcpy(hs.data(), d_start, R * arr * 12, cudaMemcpyDeviceToHost));
    CUDA_CHECK(cudaMemcpy(he.data(), d_end, R * arr * 12, cudaMemcpyDeviceToHost));
    CUDA_CHECK(cudaMemcpy(te.data(), d_te, R * arr * 4, cudaMemcpyDeviceToHost));
    CUDA_CHECK(cudaMemcpy(vd.data(), d_vd, R * 8, cudaMemcpyDeviceToHost));
    int num_points = 0;
    for (size_t r = 0; r < R; r++) { idx[r] = num_points; num_points += nh[r]; }
    std::vector<float3> cs(num_points), ce(num_points);
    std::vector<float> cte(num_points);
    for (size_t r = 0; r < R; r++)
        for (int j = 0; j < nh[r]; j++) {
            cs[idx[r] + j] = hs[r * arr + j]; ce[idx[r] + j] = he[r * arr + j]; cte[idx[r] + j] = te[r * arr + j];
        }
    const size_t N = (size_t)num_points * NUM_SAMPLES_PER_SEGMENT;
    float3* d_cs = dmalloc<float3>(num_points + 1);
    float3* d_ce = dmalloc<float3>(num_points + 1);
    int* d_idx = dmalloc<int>(R);
    float* d_t_vals = dmalloc<float>(N + 1);
    float* d_samples = dmalloc<float>(N * 5 + 1);
    float* d_radiance = dmalloc<float>(N * 4 + 1);
    float* d_pixels = dmalloc<float>(R * 3);
    CUDA_CHECK(cudaMemcpy(d_cs, cs.data(), (size_t)num_points * 12, cudaMemcpyHostToDevice));
    CUDA_CHECK(cudaMemcpy(d_ce, ce.data(), (size_t)num_points * 12, cudaMemcpyHostToDevice));
    CUDA_CHECK(cudaMemcpy(d_idx, idx.data(), R * 4, cudaMemcpyHostToDevice));
    launchSampler(d_cs, d_ce, d_vd, d_t_vals, d_samples, (int)R, G, d_nh, d_idx, SAMPLING_REGULAR, inference_stream);  // :704-712
    CUDA_CHECK(cudaStreamSynchronize(inference_stream));
    // stand-in radiance, a fixed function of the sample index (no tcnn in this test)
    std::vector<float> rad(N * 4);
    for (size_t i = 0; i < N; i++) {
        rad[4 * i] = (float)((i * 7) % 11) / 11.0f; rad[4 * i + 1] = (float)((i * 3) % 5) / 5.0f;
        rad[4 * i + 2] = (float)(i % 13) / 13.0f;  rad[4 * i + 3] = (float)((i * 5) % 17);
    }
    CUDA_CHECK(cudaMemcpy(d_radiance, rad.data(), N * 16, cudaMemcpyHostToDevice));
    launch_volrender_cuda(d_samples, d_radiance, d_nh, d_idx, d_t_vals, (int)R, NUM_SAMPLES_PER_SEGMENT, d_pixels);     // :737-746
    std::vector<__half> lg(R * 3);
    for (size_t i = 0; i < R * 3; i++) lg[i] = __float2half((float)((i % 9) - 4) / 8.0f);
    __half* d_lg = dmalloc<__half>(R * 3);
    __half* d_rg = dmalloc<__half>(N * 4 + 1);
    CUDA_CHECK(cudaMemcpy(d_lg, lg.data(), R * 6, cudaMemcpyHostToDevice));
    launch_volrender_backward_cuda(nullptr, d_lg, d_radiance, d_t_vals, d_nh, d_idx, (int)R, NUM_SAMPLES_PER_SEGMENT, d_rg);  // :767-777
    CUDA_CHECK(cudaDeviceSynchronize());
    if (rtxnerf_last_status() || rtxnerf_volrender_status()) return 4;

    std::vector<float> samples(N * 5), tv(N), px(R * 3);
    std::vector<unsigned short> rg(N * 4), lgb(R * 3);
    CUDA_CHECK(cudaMemcpy(samples.data(), d_samples, N * 20, cudaMemcpyDeviceToHost));
    CUDA_CHECK(cudaMemcpy(tv.data(), d_t_vals, N * 4, cudaMemcpyDeviceToHost));
    CUDA_CHECK(cudaMemcpy(px.data(), d_pixels, R * 12, cudaMemcpyDeviceToHost));
    CUDA_CHECK(cudaMemcpy(rg.data(), d_rg, N * 8, cudaMemcpyDeviceToHost));
    CUDA_CHECK(cudaMemcpy(lgb.data(), d_lg, R * 6, cudaMemcpyDeviceToHost));
    FILE* f = fopen(argv[5], "wb");
    if (!f) return 5;
    dump(f, nh); dump(f, idx); dump(f, vd); dump(f, cs); dump(f, ce); dump(f, cte);
    dump(f, samples); dump(f, tv); dump(f, rad); dump(f, px); dump(f, lgb); dump(f, rg);
    fclose(f);
    delete rtx_dataholder;
    printf("compat_driver ok: %zu rays, %d hits, %zu samples\n", R, num_points, N);
    return 0;
}
