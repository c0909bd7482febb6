/*
 * rtxnerf.h — C ABI of the B200-native rtx_nerf per-ray hot path (librtxnerf_b200.so).
 *
 * Drop-in boundary for owensgroup/rtx_nerf's sampler/, vol_render/ and rtx/ entry points.
 * Every entry point cites the reference interface it replaces (paths relative to the
 * reference checkout).  Conventions, unlike the reference's void launchers
 * (SURVEY §8b "Errors"):
 *   - every function returns int: 0 = RTXNERF_OK, >0 = a cudaError_t value,
 *     <0 = an RTXNERF_E* argument error; nothing throws, exits or prints;
 *   - all buffers are CALLER-OWNED raw device pointers unless a name ends in _host;
 *   - `stream` is a cudaStream_t passed as void* (NULL = legacy default stream, which is
 *     what vol_render.cu:155,179 launch on);
 *   - launches are asynchronous on `stream`; nothing synchronises unless stated.
 * The C++ shim headers in rtx_nerf_b200/compat/ re-expose the reference's exact C++
 * signatures on top of this ABI so main.cu links unchanged.
 */
#ifndef RTXNERF_H
#define RTXNERF_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RTXNERF_ABI_VERSION 1

enum {
    RTXNERF_OK = 0,
    RTXNERF_EINVAL = -1,   /* null pointer / bad size / a buffer that breaks the vector alignment its layout implies
                              (16 bytes for [N][4] fp32 rows and sample rows, 8 for float2 / four-fp16 rows) */
    RTXNERF_EGRID = -2,    /* grid resolution not a power of two in [2, 1024] */
    RTXNERF_ECAPACITY = -3,/* caller-provided capacity too small */
    RTXNERF_ENOMEM = -4,
    RTXNERF_ECONFIG = -5   /* unsupported network / encoding configuration */
};

int rtxnerf_abi_version(void);
/* Number of CUDA kernels this library has launched in this process (all entry points). */
unsigned long long rtxnerf_kernel_launches(void);
/* Static string for a code returned by this library (CUDA codes map to cudaGetErrorString). */
const char* rtxnerf_error_string(int code);

/* ------------------------------------------------------------------------------------
 * Occupancy grid — replaces RTXDataHolder::initAccelerationStructure
 * (rtx/include/rtxFunctions.h:89, rtx/src/rtxFunctions.cpp:293-351) and make_grid
 * (main.cu:154-174).  A cell p = (x*G+y)*G+z is "occupied" iff it would be an AABB
 * primitive in the reference's GAS.  Bounds are make_grid's: min = -1 + i*(2/G).
 * ---------------------------------------------------------------------------------- */
typedef struct rtxnerf_grid rtxnerf_grid;

/* occ_bits_host: G^3 bits, bit (p & 31) of word (p >> 5).  Synchronous (uploads). */
int rtxnerf_grid_create(const uint32_t* occ_bits_host, int G, rtxnerf_grid** out);
/* From the reference's own primitive list: n AABBs {minX,minY,minZ,maxX,maxY,maxZ} on the
 * host (std::vector<OptixAabb>, main.cu:396-399); each must be a make_grid(G) cell. */
int rtxnerf_grid_create_from_aabbs(const float* aabb6_host, int n_prims, int G, rtxnerf_grid** out);
/* Re-upload the occupancy bits (device-side rebuild of the skip hierarchy), async. */
int rtxnerf_grid_update(rtxnerf_grid* g, const uint32_t* occ_bits_host, void* stream);
int rtxnerf_grid_destroy(rtxnerf_grid* g);
int rtxnerf_grid_resolution(const rtxnerf_grid* g);
/* Device AABB array of all G^3 make_grid cells (what initAccelerationStructure returns and
 * main.cu:399,492 stores in Params.aabb); built lazily, owned by the grid. */
int rtxnerf_grid_device_aabbs(rtxnerf_grid* g, float** d_aabb6_out);

/* Device-side rebuild from densities (the reference has only the static make_grid
 * and its "TODO: make this a CUDA kernel", main.cu:153).
 *   rtxnerf_grid_cell_samples: [G^3][5] fp32 rows (x,y,z = cell centre, theta = phi = 0) in
 *     primitive order — the rows to run the network on;
 *   rtxnerf_grid_update_from_density: sigma of cell p at d_sigma[p*sigma_stride] (stride 4 on the
 *     network's [n][4] radiance buffer, +3 offset applied by the caller);
 *     d_ema != NULL: ema = max(ema*decay, sigma) (instant-ngp's rule) else ema = sigma;
 *     cell occupied iff ema > threshold.  Asynchronous; the next march sees the new bits.
 *   rtxnerf_grid_read_bits: synchronous copy of the bitfield to the host. */
int rtxnerf_grid_cell_samples(int G, float* d_rows5, void* stream);
int rtxnerf_grid_update_from_density(rtxnerf_grid* g, const float* d_sigma, int sigma_stride, float decay,
                                     float* d_ema, float threshold, void* stream);
int rtxnerf_grid_read_bits(const rtxnerf_grid* g, uint32_t* bits_host);

/* Headless image output (stb_image_write is vendored but unused, main.cu:19-21):
 * fp32 RGB [height][width][3] on the HOST, clamped to [0,1] -> binary PPM (P6). */
int rtxnerf_write_ppm(const char* path, const float* pixels_host, int width, int height);

/* ------------------------------------------------------------------------------------
 * Ray march — replaces optixLaunch(pipeline_ray_march, ...) (main.cu:506-508) and the
 * device programs rtx/src/optixPrograms.cu:43-248.
 * ---------------------------------------------------------------------------------- */

/* Mirror of `struct Params` (rtx/include/params.h:14-42): same field order, OptiX types
 * replaced by plain ones (OptixTraversableHandle -> uint64, OptixAabb* -> float*,
 * float3 -> float[3], float2* / float3* -> float*).  Same size and offsets on LP64. */
typedef struct rtxnerf_params {
    uint64_t handle;            /* (uint64_t)(rtxnerf_grid*) — OptixTraversableHandle slot */
    float* look_at;             /* device, 4x4 row-major pose */
    int num_primitives;
    int intersection_arr_size;  /* slots per ray (3*G at main.cu:486) */
    float* aabb;                /* unused by the new kernels (grid handle carries geometry) */
    float* start_points;        /* [W*H][arr][3] */
    float* end_points;          /* [W*H][arr][3] */
    float* ray_origins;         /* [W*H][3] */
    float* t_start;             /* [W*H][arr] */
    float* t_end;               /* [W*H][arr] */
    int* num_hits;              /* [W*H] */
    float* viewing_direction;   /* [W*H][2] (theta, phi) */
    float focal_length;
    float aspect_ratio;
    float delta[3];
    float min_point[3];
    float max_point[3];
    unsigned int width, height;
} rtxnerf_params;

/* Dense layout exactly as the reference writes it: hit j of ray r at slot
 * r*intersection_arr_size + j (optixPrograms.cu:184).  Unlike the reference (Q11) a ray
 * stops after intersection_arr_size hits.  `params_host` is read on the host. */
int rtxnerf_ray_march_dense(const rtxnerf_params* params_host, void* stream);

/* Camera block for the CSR entry points (the Params scalars + pose). */
typedef struct rtxnerf_camera {
    const float* look_at;   /* device, 16 floats row-major */
    float focal_length;
    float aspect_ratio;
    unsigned int width, height;
} rtxnerf_camera;

/* CSR hit records of rays [ray_begin, ray_begin + n_rays) of one view; all arrays are
 * indexed by LOCAL ray (ray - ray_begin) / local hit slot.  Replaces the dense buffers +
 * thrust reduce/exclusive_scan + host gather of main.cu:517-543,631-676. */
typedef struct rtxnerf_hits {
    int* num_hits;        /* [n_rays]        optixPrograms.cu:241 */
    int* indices;         /* [n_rays]        exclusive scan of num_hits (main.cu:634-637) */
    float* view_dirs;     /* [n_rays][2]     optixPrograms.cu:71-73 */
    float* ray_origins;   /* [n_rays][3]     optixPrograms.cu:75-82 (may be NULL) */
    float* t_start;       /* [capacity]      (may be NULL) */
    float* t_end;         /* [capacity]      (may be NULL) */
    float* start_points;  /* [capacity][3] */
    float* end_points;    /* [capacity][3] */
    int* hit_cells;       /* [capacity]      primitive index (x*G+y)*G+z (may be NULL) */
    long long* total_hits;/* [1] device: sum of num_hits */
    long long capacity;   /* hit slots available in the arrays above */
} rtxnerf_hits;

/* Bytes of device scratch rtxnerf_ray_march_csr needs for n_rays (any max_hits). */
size_t rtxnerf_ray_march_workspace_bytes(long long n_rays);
/* Bytes of scratch that also let a march with this max_hits (33..1024) run as ONE traversal:
 * the base size plus a staging slab for the hit records of the resident CTAs (independent of
 * n_rays beyond 148 x 8 x 128 rays).  For other max_hits it equals the base size. */
size_t rtxnerf_ray_march_workspace_bytes_for(long long n_rays, int max_hits);

/* All on `stream`.  max_hits <= 32: ONE kernel (march once, hits parked in shared memory,
 * decoupled look-back for the CSR base, coalesced copy-out) after a memset of the status
 * words.  33..1024 with a workspace of ..._bytes_for(n_rays, max_hits): the same kernel as a
 * persistent grid, hits parked in the workspace.  Otherwise count -> scan -> fill (every ray
 * walked twice).  max_hits caps hits per ray (H_max).  Hits
 * beyond `capacity` are dropped and *total_hits still reports the true total, so the
 * caller can detect overflow; rays are never split. */
int rtxnerf_ray_march_csr(const rtxnerf_grid* grid, const rtxnerf_camera* cam,
                          long long ray_begin, long long n_rays, int max_hits,
                          const rtxnerf_hits* hits, void* workspace, size_t workspace_bytes,
                          void* stream);
/* The same march that ALSO writes every hit's entry and exit point as one 32-byte record,
 * d_segments8[slot] = (sx, sy, sz, ex, ey, ez, 0, 0), [capacity][8] floats, 32-byte aligned.  The lanes of a warp
 * read the hits of 32 different rays, whose lists are ~200 bytes apart: from the two [capacity][3] arrays that is
 * six 4-byte loads of 32 different sectors each per sample, from the record two 16-byte loads of one sector per
 * lane.  Consumer: rtxnerf_render_rays_fused_seg.  hits->start_points and hits->end_points may both be NULL here
 * (the records alone are written).  (Repo-native layout; the reference keeps the two arrays,
 * optixPrograms.cu:194-211.) */
int rtxnerf_ray_march_csr_seg(const rtxnerf_grid* grid, const rtxnerf_camera* cam,
                              long long ray_begin, long long n_rays, int max_hits,
                              const rtxnerf_hits* hits, float* d_segments8, void* workspace,
                              size_t workspace_bytes, void* stream);

/* The same march with NO device allocation and no caller-provided workspace.  No counterpart in the
 * reference: added for callers that cannot call cudaMalloc (code under
 * CUDA graph capture, tight serving loops, harnesses that forbid it).  The scratch of rtxnerf_ray_march_csr
 * lives in a file-scope device array of the library (~321 KB that every GPU which loads the library holds until
 * exit), sized for at most rtxnerf_ray_march_csr_static_max_rays() = 2^21 rays per call (a 1920x1080 view
 * fits); more rays are refused with RTXNERF_ECAPACITY.  max_hits <= 32 runs the one-pass kernel, larger caps run
 * count -> scan -> fill.  Outputs are identical to rtxnerf_ray_march_csr's.  Every call on a device shares the
 * array and finds the previous call's contents in it: the march itself clears, on `stream`, the words that must
 * start at zero, and calls that OVERLAP on the GPU (different streams without a dependency) race on it —
 * serialise them.  Calls cudaMalloc / cudaFree nowhere. */
long long rtxnerf_ray_march_csr_static_max_rays(void);
int rtxnerf_ray_march_csr_static(const rtxnerf_grid* grid, const rtxnerf_camera* cam,
                                 long long ray_begin, long long n_rays, int max_hits,
                                 const rtxnerf_hits* hits, void* stream);

/* ------------------------------------------------------------------------------------
 * Sampler — replaces launchSampler (sampler/sampler.h:19-30, sampler/sampler.cu:105-131).
 * samples_per_segment is the reference's NUM_SAMPLES_PER_SEGMENT (sampler.h:4, 32) made a
 * run-time value.  sample_type: 0 REGULAR, 1 STRATIFIED_JITTERING, 2 UNIFORM (sampler.h:5-9).
 * ---------------------------------------------------------------------------------- */
int rtxnerf_launch_sampler(const float* d_start_points, const float* d_end_points,
                           const float* d_view_dirs, float* d_t_vals, float* d_sampled_points,
                           int batch_size, int grid_res, const int* d_num_hits,
                           const int* d_indices, int sample_type, int samples_per_segment,
                           void* stream);

/* Same, with counter-based random draws for the STRATIFIED_JITTERING / UNIFORM modes (SURVEY 8
 * f4): seed != 0 replaces the reference's by-value default-seeded thrust::minstd_rand
 * (sampler.cu:25,117 — every thread draws the identical sequence and every launch repeats it)
 * by u = splitmix64(seed, global sample index) >> 40 scaled to [0,1), t = fma(u, b - a, a);
 * seed == 0 is rtxnerf_launch_sampler.  REGULAR mode ignores the seed. */
int rtxnerf_launch_sampler_seeded(const float* d_start_points, const float* d_end_points,
                                  const float* d_view_dirs, float* d_t_vals, float* d_sampled_points,
                                  int batch_size, int grid_res, const int* d_num_hits,
                                  const int* d_indices, int sample_type, int samples_per_segment,
                                  unsigned long long seed, void* stream);

/* ------------------------------------------------------------------------------------
 * Volume rendering — replaces launch_volrender_cuda / launch_volrender_backward_cuda
 * (vol_render/vol_render.h:5-25, vol_render/vol_render.cu:144-190).  __half buffers are
 * passed as uint16_t* (IEEE binary16 bits).  network_inputs / loss_values are accepted
 * and ignored, as in the reference.
 * ---------------------------------------------------------------------------------- */
int rtxnerf_launch_volrender(const float* network_inputs, const float* network_outputs,
                             const int* num_hits, const int* indices, const float* ray_hit,
                             int batch_size, int num_samples_per_hit, float* pixels, void* stream);

int rtxnerf_launch_volrender_backward(const float* loss_values, const uint16_t* loss_gradients,
                                      const float* sampled_points_radiance, const float* t_hit,
                                      const int* num_hits, const int* indices, int batch_size,
                                      int num_samples_per_hit, uint16_t* radiance_gradients,
                                      void* stream);

/* The same with an explicit gradient mode.  RTXNERF_BACKWARD_REFERENCE (0, what the entry point above runs) is
 * the reference's own formula, vol_render.cu:118-135, kept bit for bit although it is NOT the gradient of the
 * forward pass.  RTXNERF_BACKWARD_ANALYTIC (1) is the exact gradient of vol_render.cu:55-67
 *   dL/dc_k = g w_k,   dL/dsigma_k = delta_k [exp(-tau_k - delta_k sigma_k) (g.c_k) - sum_{j>=k} w_j (g.c_j)],
 * for callers that want the training step to descend the loss it reports.  Same buffers and layouts. */
enum { RTXNERF_BACKWARD_REFERENCE = 0, RTXNERF_BACKWARD_ANALYTIC = 1 };
int rtxnerf_launch_volrender_backward_mode(const float* loss_values, const uint16_t* loss_gradients,
                                           const float* sampled_points_radiance, const float* t_hit,
                                           const int* num_hits, const int* indices, int batch_size,
                                           int num_samples_per_hit, uint16_t* radiance_gradients,
                                           int backward_mode, void* stream);

/* Config-1 stand-in for the network (BASELINE.json configs[0]): every sample takes the
 * fixed (r,g,b,sigma) of the cell its segment lies in.  n_hits is read from the device
 * scalar d_total_hits; at most capacity_hits segments are processed. */
int rtxnerf_cell_radiance(const int* d_hit_cells, const long long* d_total_hits,
                          long long capacity_hits, int samples_per_segment,
                          const float* d_cell_rgbs4, float* d_radiance, void* stream);

/* ------------------------------------------------------------------------------------
 * Network — replaces tiny-cuda-nn's encoding + FullyFusedMLP forward as called at
 * main.cu:715-728 (network->forward + output slice + convertHalfToFloat).  tiny-cuda-nn
 * @ d0522b4 is absent from the reference archive: the algorithms are restated from their
 * published descriptions (PARITY UNPINNED).
 * ---------------------------------------------------------------------------------- */
#define RTXNERF_MAX_LEVELS 16

/* Multiresolution hash encoding (instant-ngp) of xyz in [-1,1]^3.  Per-level geometry is
 * explicit so that host oracle and kernel share it bit for bit:
 *   pos = scale*x01 + 0.5 ; level l is dense (x + y*res + z*res^2) when res^3 <= size,
 *   else hashed (x ^ y*2654435761 ^ z*805459861) mod size.  F = n_features = 2. */
typedef struct rtxnerf_hashgrid {
    const uint16_t* table;               /* device fp16 [sum(size)][2] */
    int n_levels;                        /* 1..16 */
    int n_features;                      /* 2 */
    float scale[RTXNERF_MAX_LEVELS];
    uint32_t res[RTXNERF_MAX_LEVELS];
    uint32_t size[RTXNERF_MAX_LEVELS];   /* entries in the level */
    uint32_t offset[RTXNERF_MAX_LEVELS]; /* first entry of the level */
} rtxnerf_hashgrid;

/* Bias-free MLP: fp16 weights row-major [out][in], layers concatenated
 * [width][in_dim], (n_hidden-1) x [width][width], [16][width]; ReLU hidden, output padded
 * to 16 (main.cu:715 padded_output_width), out_activation 0 = none, 1 = sigmoid (main.cu:63). */
typedef struct rtxnerf_mlp {
    const uint16_t* weights;             /* device fp16 */
    int in_dim;                          /* 32 (= 16 levels x 2 features) */
    int width;                           /* 64 */
    int n_hidden;                        /* 2 */
    int out_activation;
} rtxnerf_mlp;

enum { RTXNERF_IMPL_TCGEN05 = 0, RTXNERF_IMPL_SIMT = 1 };

/* The number of rows processed is n = min(n_max, d_count ? *d_count * count_scale : n_max):
 * d_count is the device hit total of the march, count_scale the samples per segment. */

/* samples rows of `sample_stride` floats (x,y,z first) -> encoded fp16 [n][n_levels*2]. */
int rtxnerf_hashgrid_encode(const rtxnerf_hashgrid* enc, const float* d_samples, int sample_stride,
                            const long long* d_count, long long count_scale, long long n_max,
                            uint16_t* d_encoded, void* stream);

/* enc != NULL: input = fp32 sample rows [n][input_stride] (the sampler's [n][5] buffer,
 * main.cu:716 GPUMatrix<float>(5,N)); enc == NULL: input = pre-encoded fp16 [n][in_dim].
 * out16 (optional): fp16 [n][16] = GPUMatrix<half>(16,N) column-major (main.cu:720-721).
 * radiance4 (optional): fp32 [n][4] = rows 0..3 converted to float (main.cu:723-728, the
 * intended gather of Q3, fused into the epilogue).  impl selects the tcgen05 kernel (product)
 * or the CUDA-core check kernel. */
int rtxnerf_network_forward(const rtxnerf_hashgrid* enc, const rtxnerf_mlp* mlp, const void* d_input,
                            int input_stride, const long long* d_count, long long count_scale,
                            long long n_max, uint16_t* d_out16, float* d_radiance4, int impl,
                            void* stream);

/* Ray-coherent schedule of the same function over the sampler's CSR sample buffer
 * (sample (j,i) of ray x at row (indices[x]+j)*S+i, sampler/sampler.cu:60-64): a warp covers 32
 * neighbouring rays at one sample slot, so its hash-table gathers coalesce.  Results are
 * identical to rtxnerf_network_forward on the same rows; rows not covered by
 * (num_hits, indices) are not written.  This is what the render pipeline calls.
 * d_encoded_out (optional): fp16 [n][32] encoded features, kept by a TRAINING forward
 * (network->forward(..., prepare_input_gradients), main.cu:721) for the backward pass. */
int rtxnerf_network_forward_rays(const rtxnerf_hashgrid* enc, const rtxnerf_mlp* mlp,
                                 const float* d_samples, int sample_stride, const int* d_num_hits,
                                 const int* d_indices, int n_rays, int samples_per_segment,
                                 uint16_t* d_out16, float* d_radiance4, uint16_t* d_encoded_out,
                                 void* stream);

/* The same forward without a sample buffer: sample (j,i) of ray x is rebuilt from the packed hit records of
 * rtxnerf_ray_march_csr_seg / rtxnerf_batch_assemble_seg with the sampler's own arithmetic
 * (sampler/sampler.cu:45-59, REGULAR sampling); row order and results as rtxnerf_launch_sampler followed by
 * rtxnerf_network_forward_rays.  A training step that uses this and rtxnerf_hashgrid_backward_seg launches no
 * sampler and keeps no [N][5] sample buffer (main.cu:704-721). */
int rtxnerf_network_forward_seg(const rtxnerf_hashgrid* enc, const rtxnerf_mlp* mlp,
                                const float* d_segments8, const int* d_num_hits, const int* d_indices,
                                int n_rays, int samples_per_segment, uint16_t* d_out16,
                                float* d_radiance4, uint16_t* d_encoded_out, void* stream);

/* ------------------------------------------------------------------------------------
 * The reference's literal network (main.cu:47-68): Composite[Frequency(xyz, 10 frequencies) +
 * Frequency(theta/phi, tiny-cuda-nn's default 12)] padded with ones to 112 inputs ->
 * FullyFusedMLP 8 x 128 ReLU -> 16 outputs (sigmoid), bias-free fp16.  Feature order:
 * out[d*2F + 2k] = sin(pi 2^k x_d), out[d*2F + 2k + 1] = cos(pi 2^k x_d).  Weights: fp16
 * row-major [128][112], 7 x [128][128], [16][128] concatenated (rtxnerf_freqnet_param_count()
 * values); weights_packed is the same data in the tensor-core operand layout, produced once by
 * rtxnerf_freqnet_pack (repeat after every optimizer step).  PARITY UNPINNED (tiny-cuda-nn
 * absent), see DESIGN.md.
 * ---------------------------------------------------------------------------------- */
typedef struct rtxnerf_freqnet {
    const uint16_t* weights_packed;  /* device, 16-byte aligned */
    int n_freq_xyz;                  /* 10 */
    int n_freq_dir;                  /* 12 */
    int width;                       /* 128 */
    int n_hidden;                    /* 8 */
    int out_activation;              /* 0 none, 1 sigmoid */
} rtxnerf_freqnet;

size_t rtxnerf_freqnet_param_count(void);    /* fp16 values in `weights` */
size_t rtxnerf_freqnet_packed_count(void);   /* fp16 values in `weights_packed` (layer 0 padded to 128 inputs) */
int rtxnerf_freqnet_pack(const uint16_t* d_weights, uint16_t* d_packed, void* stream);
/* [n][sample_stride >= 5] fp32 rows (x,y,z,theta,phi) -> fp16 [n][112] */
int rtxnerf_freqnet_encode(const float* d_samples, int sample_stride, long long n, uint16_t* d_encoded,
                           void* stream);
/* Rows as rtxnerf_network_forward: n = min(n_max, d_count ? *d_count * count_scale : n_max);
 * d_out16 fp16 [n][16] and/or d_radiance4 fp32 [n][4]. */
int rtxnerf_freqnet_forward(const rtxnerf_freqnet* net, const float* d_samples, int sample_stride,
                            const long long* d_count, long long count_scale, long long n_max,
                            uint16_t* d_out16, float* d_radiance4, void* stream);

/* Backward of that network (network->backward, main.cu:781): fp32 gradients of the row-major weights (the
 * layout of rtxnerf_freqnet_pack's input), ACCUMULATED into d_grad_weights (caller-zeroed; rtxnerf_adam_step
 * re-zeroes).  The Frequency encoding has no parameters, so nothing else is produced.  d_dl_dradiance: fp16
 * [n][4] as launch_volrender_backward_cuda writes it.  The forward is recomputed from d_samples per 128-row
 * tile.  Rows as rtxnerf_freqnet_forward.  PARITY UNPINNED (tiny-cuda-nn absent). */
int rtxnerf_freqnet_backward(const uint16_t* d_weights, int out_activation, const float* d_samples,
                             int sample_stride, const uint16_t* d_dl_dradiance, const long long* d_count,
                             long long count_scale, long long n_max, float* d_grad_weights, void* stream);

/* Render path in ONE kernel: REGULAR sampling of the CSR segments (launchSampler,
 * sampler/sampler.cu:14-131), encoding + MLP forward (main.cu:715-728) and front-to-back
 * compositing (launch_volrender_cuda, vol_render/vol_render.cu:19-73) without the samples /
 * t_vals / radiance buffers ever reaching HBM.  Same arithmetic per sample as the three staged
 * entry points; the per-ray accumulation runs in the reference's sequential order.
 * sample_type must be 0 (REGULAR).  pixels: [n_rays][3] fp32. */
int rtxnerf_render_rays_fused(const rtxnerf_hashgrid* enc, const rtxnerf_mlp* mlp,
                              const float* d_start_points, const float* d_end_points,
                              const int* d_num_hits, const int* d_indices, int n_rays,
                              int samples_per_segment, int sample_type, float* d_pixels, void* stream);
/* rtxnerf_render_rays_fused over the packed hit records of rtxnerf_ray_march_csr_seg: same samples, same pixels. */
int rtxnerf_render_rays_fused_seg(const rtxnerf_hashgrid* enc, const rtxnerf_mlp* mlp,
                                  const float* d_segments8, const int* d_num_hits, const int* d_indices,
                                  int n_rays, int samples_per_segment, int sample_type, float* d_pixels,
                                  void* stream);

/* ------------------------------------------------------------------------------------
 * Training step — replaces network->backward (main.cu:781), loss->evaluate + reduce_sum
 * (main.cu:755-760) and optimizer->step (main.cu:787) of tiny-cuda-nn (absent from the
 * archive; restated, PARITY UNPINNED).  Gradients are fp32 and ACCUMULATE into caller-owned,
 * caller-zeroed buffers (rtxnerf_adam_step re-zeroes them), so that one NCCL all-reduce of
 * [grad_table | grad_weights] is the only exchange of a data-parallel step.
 * ---------------------------------------------------------------------------------- */

/* d_dl_dradiance: fp16 [n][4], what launch_volrender_backward_cuda wrote (the wiring the driver
 * intends at main.cu:765-780, Q5).  d_encoded: the training forward's features.  d_dencoded:
 * fp16 [n][32] scratch (dL/d features, written then scattered).  Rows: n = min(n_max,
 * *d_count * samples_per_segment) in CSR order.  grad_table: fp32 [entries][2]; grad_weights:
 * fp32, same layout as rtxnerf_mlp.weights. */
int rtxnerf_network_backward_rays(const rtxnerf_hashgrid* enc, const rtxnerf_mlp* mlp,
                                  const float* d_samples, int sample_stride, const uint16_t* d_encoded,
                                  const uint16_t* d_dl_dradiance, const int* d_num_hits,
                                  const int* d_indices, int n_rays, int samples_per_segment,
                                  const long long* d_count, long long n_max, uint16_t* d_dencoded,
                                  float* d_grad_table, float* d_grad_weights, void* stream);

/* The same two kernels as separate calls (rtxnerf_network_backward_rays = the first, then the second on the same
 * stream).  d_count * count_scale (capped at n_max) is the number of sample rows; a data-parallel caller
 * all-reduces the MLP weight gradients after the first call, under the scatter of the second. */
int rtxnerf_mlp_backward(const rtxnerf_mlp* mlp, const uint16_t* d_encoded, const uint16_t* d_dl_dradiance,
                         const long long* d_count, long long count_scale, long long n_max, uint16_t* d_dencoded,
                         float* d_grad_weights, void* stream);
int rtxnerf_hashgrid_backward_rays(const rtxnerf_hashgrid* enc, const float* d_samples, int sample_stride,
                                   const uint16_t* d_dencoded, const int* d_num_hits, const int* d_indices,
                                   int n_rays, int samples_per_segment, float* d_grad_table, void* stream);
/* rtxnerf_hashgrid_backward_rays with the positions rebuilt from the packed hit records (REGULAR sampling)
 * instead of read from the sample buffer: the backward half of rtxnerf_network_forward_seg (main.cu:781). */
int rtxnerf_hashgrid_backward_seg(const rtxnerf_hashgrid* enc, const float* d_segments8,
                                  const uint16_t* d_dencoded, const int* d_num_hits, const int* d_indices,
                                  int n_rays, int samples_per_segment, float* d_grad_table, void* stream);

/* values[i] = (pred-target)^2 / n (optional), gradients[i] = fp16(loss_scale * 2 (pred-target) / n),
 * *d_loss_sum += sum(values) (optional, caller-zeroed).  n = batch_size * 3. */
int rtxnerf_l2_loss(const float* d_pred, const float* d_target, long long n, float loss_scale,
                    float* d_values, uint16_t* d_gradients, float* d_loss_sum, void* stream);

/* Adam over n parameters: g = grads * grad_scale (1 / (loss_scale * ranks)); grads are zeroed;
 * master (fp32) and params (fp16 copy) updated; step counts from 1 (bias correction). */
int rtxnerf_adam_step(long long n, float* d_master, uint16_t* d_params, float* d_grads, float* d_m1,
                      float* d_m2, float grad_scale, float lr, float beta1, float beta2, float eps,
                      int step, void* stream);
/* The same with tiny-cuda-nn's treatment of encoding parameters: the first n_encoding parameters (the hash
 * table in this repo's [table | weights] layout) are left untouched — weight and both moments — in a step
 * where their gradient is exactly zero, i.e. when no sample of the batch touched the entry; the remaining
 * (matrix) parameters always take the update.  n_encoding = 0 is rtxnerf_adam_step. */
int rtxnerf_adam_step_sparse(long long n, long long n_encoding, float* d_master, uint16_t* d_params,
                             float* d_grads, float* d_m1, float* d_m2, float grad_scale, float lr, float beta1,
                             float beta2, float eps, int step, void* stream);

/* ------------------------------------------------------------------------------------
 * Seeded ray shuffle + batch assembly — replaces the front end of the reference's training
 * loop (main.cu:612-676): std::random_shuffle of the host ray payloads (:615), the host loops
 * collecting num_hits / gt pixel / view_dir of a batch (:622-626), their H2D copies (:627-629),
 * thrust::reduce + exclusive_scan (:631-637) and the host gather + H2D of start/end points
 * (:649-676).  The march results of all views stay on the device (the "ray store", filled by
 * rtxnerf_ray_march_csr view after view); a batch is a slice of a stateless seeded permutation.
 * ---------------------------------------------------------------------------------- */
typedef struct rtxnerf_ray_store {
    long long n_rays;             /* n_views * rays_per_view */
    long long rays_per_view;
    const long long* view_base;   /* device [n_views]: first hit slot of each view in the record arrays */
    const int* num_hits;          /* [n_rays] */
    const int* indices;           /* [n_rays]  CSR index LOCAL to the ray's view (as the march wrote it) */
    const float* view_dirs;       /* [n_rays][2] */
    const float* start_points;    /* [total hits][3] */
    const float* end_points;      /* [total hits][3] */
    const float* t_start;         /* [total hits] (may be NULL) */
    const float* t_end;           /* [total hits] (may be NULL) */
    const float* pixels_gt;       /* [n_rays][3] ground-truth colours (may be NULL) */
} rtxnerf_ray_store;

/* d_out[k] = ray id at position first + k of the shuffled order of n rays, 0 <= k < count.
 * The permutation is a keyed bijection (4-round Feistel network, cycle-walked) over groups of
 * `group` consecutive rays (n % group == 0): group = 1 shuffles single rays as the reference
 * does, group = 32 keeps 32 adjacent pixels together.  A new epoch = a new seed. */
int rtxnerf_ray_permutation(long long n, unsigned long long seed, int group, long long first,
                            long long count, long long* d_out, void* stream);

size_t rtxnerf_batch_workspace_bytes(long long batch_rays);

/* Batch = positions [first, first + batch_rays) of the shuffled order (whole batches only,
 * first + batch_rays <= n_rays; the reference slices past the end, Q6).  Writes into `batch`:
 * num_hits, indices (exclusive scan), view_dirs, start_points, end_points, t_start / t_end
 * (when present on both sides), *total_hits; ray_origins and hit_cells are not touched.  Hits
 * beyond batch->capacity are dropped (total still true).  d_gt_pixels [batch_rays][3] and
 * d_ray_ids [batch_rays] are optional.  Three kernels on `stream`. */
int rtxnerf_batch_assemble(const rtxnerf_ray_store* store, unsigned long long seed, int group,
                           long long first, long long batch_rays, const rtxnerf_hits* batch,
                           float* d_gt_pixels, long long* d_ray_ids, void* workspace,
                           size_t workspace_bytes, void* stream);
/* rtxnerf_batch_assemble that also writes the batch's hits as the packed 32-byte records of
 * rtxnerf_ray_march_csr_seg (d_segments8 [batch->capacity][8], 32-byte aligned); batch->start_points and
 * batch->end_points may both be NULL (records only). */
int rtxnerf_batch_assemble_seg(const rtxnerf_ray_store* store, unsigned long long seed, int group,
                               long long first, long long batch_rays, const rtxnerf_hits* batch,
                               float* d_segments8, float* d_gt_pixels, long long* d_ray_ids,
                               void* workspace, size_t workspace_bytes, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* RTXNERF_H */
