/*
 * rtxnerf_oracle.h — CPU restatement of the rtx_nerf per-ray hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is part of the product:
 * only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * `--impl reference` legs may load it, and only as the checker / the timed CPU
 * baseline.  The product path (rtx_nerf_b200/) never links or imports it.
 *
 * Every function cites the reference file:line (relative to the reference
 * checkout) whose arithmetic it restates.  Floating-point contraction follows
 * the reference's own PTX (nvcc -fmad=true): each place where nvcc emits
 * fma.rn.f32 is an explicit fmaf() here, and this file is compiled with
 * -ffp-contract=off so nothing else fuses.
 *
 * Parity pin status (see DESIGN.md "Oracle"):
 *   sampler / vol_render        pinned against the reference's own kernels
 *                               (oracle/_ref host build, and sm_100a build on the GPU box)
 *   ray march (OptiX programs)  pinned against the reference's own device programs
 *                               compiled for the host over an emulated OptiX runtime
 *                               (oracle/host_shim/optix.h); OptiX itself is closed,
 *                               so closest-hit selection is a declared convention
 *   hash-grid / MLP / loss / Adam   PARITY UNPINNED (tiny-cuda-nn submodule absent)
 */
#ifndef RTXNERF_ORACLE_H
#define RTXNERF_ORACLE_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* main.cu:154-174  make_grid: dense G^3 AABB list, prim = (x*G+y)*G+z, 6 floats each
 * {minX,minY,minZ,maxX,maxY,maxZ}. */
void orc_make_grid(int G, float* aabb6);

/* optixPrograms.cu:43-82  pinhole ray for pixel (x,y): origin o, unit direction d,
 * view_dir = (theta, phi). */
void orc_raygen(const float* look_at16, float focal_length, float aspect_ratio,
                unsigned width, unsigned height, unsigned x, unsigned y,
                float o[3], float d[3], float view_dir[2]);

/* Output block shared by both marchers.  Slot of hit j of ray r is
 *   base(r) + j,  base(r) = indices ? indices[r] : r * arr_size      (optixPrograms.cu:184)
 * Any pointer except num_hits may be NULL (not written).  hit_prims receives the
 * primitive index (x*G+y)*G+z of each hit cell (the "hit-cell list"). */
typedef struct {
    float* t_start;      /* [slots]      optixPrograms.cu:188-189 */
    float* t_end;        /* [slots]      optixPrograms.cu:239     */
    float* start_points; /* [slots][3]   optixPrograms.cu:194-197,235 */
    float* end_points;   /* [slots][3]   optixPrograms.cu:209-211,236 */
    int*   hit_prims;    /* [slots]      primitive index of the hit AABB */
    int*   num_hits;     /* [R]          optixPrograms.cu:241     */
    float* view_dirs;    /* [R][2]       optixPrograms.cu:71-73   */
    float* ray_origins;  /* [R][3]       optixPrograms.cu:75-82   */
    const int* indices;  /* [R] or NULL  CSR offsets (main.cu:634-637) */
    int arr_size;        /* dense stride (3*G in main.cu:486) when indices==NULL */
    int max_hits;        /* per-ray cap (the reference has none: Q11); <=0 means arr_size */
} orc_march_out;

/* Literal OptiX semantics: every relaunch tests ALL n_prims AABBs with the slab
 * predicate (optixPrograms.cu:132-169), accepts t in [0,102], takes the smallest t
 * (ties: lowest primitive index — declared convention), runs closest-hit
 * (optixPrograms.cu:180-248) and relaunches from the exit point (optixPrograms.cu:99-115).
 * prim_ids maps list position -> primitive index (NULL: identity).  Rays are the
 * width*height pixels, ray = x + y*width (optixPrograms.cu:45), restricted to
 * [ray_begin, ray_end).  Returns the number of rays that stopped on the cap. */
int orc_ray_march_brute(const float* look_at16, float focal_length, float aspect_ratio,
                        unsigned width, unsigned height,
                        const float* aabb6, const int* prim_ids, int n_prims,
                        unsigned ray_begin, unsigned ray_end,
                        orc_march_out* out, int nthreads);

/* Same results, found by an exact 3-axis interval sweep (DDA) over a power-of-two
 * grid: occ[p] != 0 marks cell p = (x*G+y)*G+z as a primitive with the make_grid
 * bounds.  Returns -1 if G is not a power of two. */
int orc_ray_march_dda(const float* look_at16, float focal_length, float aspect_ratio,
                      unsigned width, unsigned height,
                      const uint8_t* occ, int G,
                      unsigned ray_begin, unsigned ray_end,
                      orc_march_out* out, int nthreads);

/* main.cu:631-637  exclusive scan of num_hits; returns the total. */
long long orc_exclusive_scan(const int* num_hits, int n, int* indices);

/* main.cu:646-676  gather dense [R][arr_size] hit records into CSR order. */
void orc_compact_hits(const int* num_hits, const int* indices, int n_rays, int arr_size,
                      const float* dense, int comps, float* compact);

/* sampler/sampler.cu:14-103 with NUM_SAMPLES_PER_SEGMENT = S (reference: 32).
 * sample_type: 0 REGULAR, 1 STRATIFIED_JITTERING, 2 UNIFORM (sampler.h:5-9).
 * samples [N][5] = x,y,z,theta,phi;  t_vals [N];  N = S * sum(num_hits). */
float orc_counter_uniform(unsigned long long seed, unsigned long long index);
void orc_sampler_seeded(const float* start_points, const float* end_points, const float* view_dirs,
                        float* t_vals, float* samples, int batch_size, const int* num_hits,
                        const int* indices, int S, int sample_type, unsigned long long seed, int nthreads);
void orc_sampler(const float* start_points, const float* end_points, const float* view_dirs,
                 float* t_vals, float* samples, int batch_size,
                 const int* num_hits, const int* indices, int S, int sample_type,
                 int nthreads);

/* Config-1 stand-in for the MLP: every sample takes the fixed (r,g,b,sigma) of the
 * cell its segment lies in.  radiance [N][4]. */
void orc_cell_radiance(const int* hit_prims, long long n_hits, int S,
                       const float* cell_rgbs4, float* radiance);

/* vol_render/vol_render.cu:19-73 */
void orc_volrender(const float* network_outputs, const int* num_hits, const int* indices,
                   const float* ray_hit, int batch_size, int S, float* pixels, int nthreads);

/* vol_render/vol_render.cu:75-143; loss_gradients and radiance_gradients are IEEE
 * binary16 bit patterns. */
void orc_volrender_backward(const uint16_t* loss_gradients, const float* radiance,
                            const float* t_hit, const int* num_hits, const int* indices,
                            int batch_size, int S, uint16_t* radiance_gradients, int nthreads);

/* ---- encoding + MLP (PARITY UNPINNED: restated from the published algorithms) ---- */

typedef struct {
    int n_levels;          /* L  */
    int n_features;        /* F (2) */
    int log2_hashmap_size; /* log2 T */
    int base_resolution;   /* N_min */
    float per_level_scale; /* b */
} orc_hashgrid_cfg;

/* Offsets (in entries, not features) of each level inside the table; returns total
 * entries.  offsets has n_levels+1 slots. */
long long orc_hashgrid_offsets(const orc_hashgrid_cfg* cfg, long long* offsets);

/* Per-level geometry: scale_l, res_l, size_l (entries) for l < n_levels. */
void orc_hashgrid_geometry(const orc_hashgrid_cfg* cfg, float* scales, uint32_t* ress, uint32_t* sizes);

/* Multiresolution hash encoding of xyz in [-1,1]^3 (samples rows, stride floats):
 * explicit per-level geometry; table fp16 [entries][F], out fp16 [N][L*F]. */
void orc_frequency_encode(const float* samples, int stride, long long n, int n_freq_xyz, int n_freq_dir,
                          int padded_width, uint16_t* out, int nthreads);
void orc_hashgrid_encode(int L, int F, const float* scales, const uint32_t* ress,
                         const uint32_t* sizes, const long long* offsets, const uint16_t* table,
                         const float* samples, int sample_stride, long long n, uint16_t* encoded,
                         int nthreads);

/* Bias-free MLP, fp16 weights row-major [out][in] per layer, fp32 accumulate, ReLU
 * hidden activations rounded to fp16, output padded to 16, optional sigmoid.
 * in fp16 [N][in_dim]; out fp16 [N][16]; radiance (optional) fp32 [N][4] = first 4. */
void orc_mlp_forward(const uint16_t* weights, int in_dim, int width, int n_hidden,
                     int out_activation /*0 none, 1 sigmoid*/,
                     const uint16_t* in, long long n, uint16_t* out16, float* radiance4,
                     int nthreads);

/* Training-step restatements (tcnn backward / L2 loss / Adam: PARITY UNPINNED).  32 -> 64 -> 64
 * -> 16 network only.  grad_w: double, layout of `weights`; grad_table: double [entries][2],
 * caller-zeroed (accumulates). */
void orc_mlp_backward(const uint16_t* weights, int out_activation, const uint16_t* enc,
                      const uint16_t* dl_dradiance, long long n, uint16_t* d_enc, double* grad_w);
void orc_freqnet_backward(const uint16_t* weights, int out_activation, const uint16_t* enc,
                          const uint16_t* dl_dradiance, long long n, double* grad_w);
void orc_hashgrid_backward(int L, const float* scales, const uint32_t* ress, const uint32_t* sizes,
                           const long long* offsets, const float* samples, int sample_stride,
                           long long n, const uint16_t* d_enc, double* grad_table);

/* fp16 helpers (round-to-nearest-even), exported for the tests. */
uint16_t orc_f2h(float f);
float    orc_h2f(uint16_t h);

#ifdef __cplusplus
}
#endif
#endif
