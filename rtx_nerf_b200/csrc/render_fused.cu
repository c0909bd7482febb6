// render_fused.cu — sampler + hash encoding + fused MLP + alpha compositing in ONE kernel, and
// the staged network forward, both on the decoupled-role kernel of rays_roles.cuh.
//
// The render path of the reference's per-ray pipeline after the march:
//   launchSampler        sampler/sampler.cu:14-131      (REGULAR mode)
//   network->forward     main.cu:715-728                (hash-grid encoding + 2x64 MLP)
//   launch_volrender     vol_render/vol_render.cu:19-73,144-164
// The staged entry points keep the reference's buffers (20 B of sample + 4 B of t + 16 B of
// radiance written and read back per sample); here a sample exists only in registers and TMEM and
// the only output is pixels[n_rays][3].
#include "rays_roles.cuh"

int vf1_render_rays_fused(const rtxnerf_hashgrid* enc, const rtxnerf_mlp* mlp, const float* d_start_points,
                          const float* d_end_points, const int* d_num_hits, const int* d_indices, int n_rays,
                          int samples_per_segment, int sample_type, float* d_pixels, void* stream);

template <bool RENDER, int EG, int LPI, bool LEAN>
static int rr_launch_t(const LeanLevels& ll, const Levels& lv, const rtxnerf_hashgrid* enc, const rtxnerf_mlp* mlp,
                       const RrArgs& a, cudaStream_t stream)
{
    static RtxPerDeviceOnce attr;
    RTX_CUDA_TRY((cudaError_t)rtx_set_max_smem(attr, rays_roles_kernel<RENDER, EG, LPI, LEAN>, RrSmem::BYTES));
    const long long groups = ((long long)a.n_rays + 31) / 32;
    // one CTA per RR_GROUPS_PER_CTA consecutive ray groups; 2 CTAs are resident per SM and the
    // hardware scheduler balances the rest
    const long long blocks = (groups + a.groups_per_cta - 1) / a.groups_per_cta;
    if (blocks > 0x7fffffffLL) return RTXNERF_EINVAL;
    rays_roles_kernel<RENDER, EG, LPI, LEAN><<<(unsigned)blocks, RR_CTA_THREADS(EG, RENDER), RrSmem::BYTES, stream>>>(
        ll, lv, reinterpret_cast<const uint32_t*>(enc->table), mlp->weights, a);
    return rtx_launch_status();
}

template <bool RENDER>
static int rr_dispatch(const LeanLevels& ll, const Levels& lv, const rtxnerf_hashgrid* enc, const rtxnerf_mlp* mlp,
                       const RrArgs& a, cudaStream_t stream)
{
    if (!ll.ok) return rr_launch_t<RENDER, 3, 1, false>(ll, lv, enc, mlp, a, stream);
    // role layout (development knob RTXNERF_RR_LAYOUT = 10 * encoder groups + levels per iteration):
    // 3 groups = 64 registers (56 with the compositor warp), 4 = 48
    static const int layout = rtx_env_int("RTXNERF_RR_LAYOUT", 32);
    switch (layout) {
    case 31: return rr_launch_t<RENDER, 3, 1, true>(ll, lv, enc, mlp, a, stream);
    case 41: return rr_launch_t<RENDER, 4, 1, true>(ll, lv, enc, mlp, a, stream);
    case 42: return rr_launch_t<RENDER, 4, 2, true>(ll, lv, enc, mlp, a, stream);
    default: return rr_launch_t<RENDER, 3, 2, true>(ll, lv, enc, mlp, a, stream);
    }
}

int rr_launch(bool render, const rtxnerf_hashgrid* enc, const rtxnerf_mlp* mlp, const Levels& lv, const RrArgs& a_in,
              cudaStream_t stream)
{
    LeanLevels ll;
    lean_levels_from(lv, &ll);
    RrArgs a = a_in;
    static const int mlp_last = rtx_env_int("RTXNERF_RR_MLP_LAST", 1);
    a.mlp_last = mlp_last;
    static const int gpc = rtx_env_int("RTXNERF_RR_GROUPS_PER_CTA", RR_GROUPS_PER_CTA);
    a.groups_per_cta = gpc > 0 ? gpc : RR_GROUPS_PER_CTA;
    static const int wait_ticks = rtx_env_int("RTXNERF_RR_WAIT_TICKS", 0);
    a.wait_ticks = wait_ticks;
    return render ? rr_dispatch<true>(ll, lv, enc, mlp, a, stream) : rr_dispatch<false>(ll, lv, enc, mlp, a, stream);
}

extern "C" int rtxnerf_render_rays_fused(const rtxnerf_hashgrid* enc, const rtxnerf_mlp* mlp,
                                         const float* d_start_points, const float* d_end_points,
                                         const int* d_num_hits, const int* d_indices, int n_rays,
                                         int samples_per_segment, int sample_type, float* d_pixels, void* stream)
{
    if (!enc || !mlp || !mlp->weights || !d_start_points || !d_end_points || !d_num_hits || !d_indices || !d_pixels)
        return RTXNERF_EINVAL;
    if (n_rays < 0 || samples_per_segment < 1) return RTXNERF_EINVAL;
    if (sample_type != 0) return RTXNERF_ECONFIG;   // REGULAR only (the reference's parity mode)
    if (mlp->in_dim != NET_IN || mlp->width != NET_W || mlp->n_hidden != 2) return RTXNERF_ECONFIG;
    if (mlp->out_activation != 0 && mlp->out_activation != 1) return RTXNERF_ECONFIG;
    static const int v1 = rtx_env_int("RTXNERF_FUSED_V1", 0);   // development: round-1 kernel for A/B runs
    if (v1)
        return vf1_render_rays_fused(enc, mlp, d_start_points, d_end_points, d_num_hits, d_indices, n_rays,
                                     samples_per_segment, sample_type, d_pixels, stream);
    Levels lv;
    int rc = levels_from(enc, &lv);
    if (rc) return rc;
    if (enc->n_levels * 2 > NET_IN) return RTXNERF_ECONFIG;
    if (n_rays == 0) return RTXNERF_OK;
    RrArgs a = {};
    a.start_points = d_start_points;
    a.end_points = d_end_points;
    a.pixels = d_pixels;
    a.num_hits = d_num_hits;
    a.indices = d_indices;
    a.n_rays = n_rays;
    a.S = samples_per_segment;
    a.out_activation = mlp->out_activation;
    return rr_launch(true, enc, mlp, lv, a, (cudaStream_t)stream);
}

// rtxnerf_render_rays_fused reading the hits as the packed 32-byte records rtxnerf_ray_march_csr_seg writes
// (d_segments8[slot] = sx, sy, sz, ex, ey, ez, 0, 0): same samples, same pixels.
extern "C" int rtxnerf_render_rays_fused_seg(const rtxnerf_hashgrid* enc, const rtxnerf_mlp* mlp,
                                             const float* d_segments8, const int* d_num_hits, const int* d_indices,
                                             int n_rays, int samples_per_segment, int sample_type, float* d_pixels,
                                             void* stream)
{
    if (!enc || !mlp || !mlp->weights || !d_segments8 || !d_num_hits || !d_indices || !d_pixels) return RTXNERF_EINVAL;
    if (((uintptr_t)d_segments8 & 31u) || n_rays < 0 || samples_per_segment < 1) return RTXNERF_EINVAL;
    if (sample_type != 0) return RTXNERF_ECONFIG;
    if (mlp->in_dim != NET_IN || mlp->width != NET_W || mlp->n_hidden != 2) return RTXNERF_ECONFIG;
    if (mlp->out_activation != 0 && mlp->out_activation != 1) return RTXNERF_ECONFIG;
    Levels lv;
    int rc = levels_from(enc, &lv);
    if (rc) return rc;
    if (enc->n_levels * 2 > NET_IN) return RTXNERF_ECONFIG;
    if (n_rays == 0) return RTXNERF_OK;
    RrArgs a = {};
    a.seg8 = reinterpret_cast<const float4*>(d_segments8);
    a.pixels = d_pixels;
    a.num_hits = d_num_hits;
    a.indices = d_indices;
    a.n_rays = n_rays;
    a.S = samples_per_segment;
    a.out_activation = mlp->out_activation;
    return rr_launch(true, enc, mlp, lv, a, (cudaStream_t)stream);
}
