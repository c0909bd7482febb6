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

template <bool RENDER, int EG>
static int rr_launch_t(const LeanLevels& ll, const Levels& lv, const rtxnerf_hashgrid* enc, const rtxnerf_mlp* mlp,
                       const RrArgs& a, cudaStream_t stream)
{
    static RtxPerDeviceOnce attr;
    RTX_CUDA_TRY((cudaError_t)rtx_set_max_smem(attr, rays_roles_kernel<RENDER, EG>, RrSmem::BYTES));
    const long long groups = ((long long)a.n_rays + 31) / 32;
    // one CTA per RR_GROUPS_PER_CTA consecutive ray groups; 2 CTAs are resident per SM and the
    // hardware scheduler balances the rest
    const long long blocks = (groups + RR_GROUPS_PER_CTA - 1) / RR_GROUPS_PER_CTA;
    if (blocks > 0x7fffffffLL) return RTXNERF_EINVAL;
    rays_roles_kernel<RENDER, EG><<<(unsigned)blocks, RR_THREADS(EG), RrSmem::BYTES, stream>>>(
        ll, lv, reinterpret_cast<const uint32_t*>(enc->table), mlp->weights, a);
    return rtx_launch_status();
}

int rr_launch(bool render, const rtxnerf_hashgrid* enc, const rtxnerf_mlp* mlp, const Levels& lv, const RrArgs& a_in,
              cudaStream_t stream)
{
    LeanLevels ll;
    lean_levels_from(lv, &ll);
    // encoder groups per CTA (development knob): 3 = 64 registers, 4 = 48, 5 = 40
    static const int eg = rtx_env_int("RTXNERF_RR_GROUPS", 3);
    static const int mlp_last = rtx_env_int("RTXNERF_RR_MLP_LAST", 1);
    RrArgs a = a_in;
    a.mlp_last = mlp_last;
    static const int wait_ticks = rtx_env_int("RTXNERF_RR_WAIT_TICKS", 0);
    a.wait_ticks = wait_ticks;
    static const int dev_skip = rtx_env_int("RTXNERF_RR_DEV_SKIP", 0);
    a.dev_skip = dev_skip;
    if (render) {
        if (eg == 4) return rr_launch_t<true, 4>(ll, lv, enc, mlp, a, stream);
        if (eg == 5) return rr_launch_t<true, 5>(ll, lv, enc, mlp, a, stream);
        return rr_launch_t<true, 3>(ll, lv, enc, mlp, a, stream);
    }
    if (eg == 4) return rr_launch_t<false, 4>(ll, lv, enc, mlp, a, stream);
    if (eg == 5) return rr_launch_t<false, 5>(ll, lv, enc, mlp, a, stream);
    return rr_launch_t<false, 3>(ll, lv, enc, mlp, a, stream);
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
