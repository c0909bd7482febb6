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

int rr_launch(bool render, const rtxnerf_hashgrid* enc, const rtxnerf_mlp* mlp, const Levels& lv, const RrArgs& a,
              cudaStream_t stream)
{
    LeanLevels ll;
    lean_levels_from(lv, &ll);
    static RtxPerDeviceOnce attr_render, attr_staged;
    if (render)
        RTX_CUDA_TRY((cudaError_t)rtx_set_max_smem(attr_render, rays_roles_kernel<true>, RrSmem::BYTES));
    else
        RTX_CUDA_TRY((cudaError_t)rtx_set_max_smem(attr_staged, rays_roles_kernel<false>, RrSmem::BYTES));
    const long long groups = ((long long)a.n_rays + 31) / 32;
    // one CTA per RR_GROUPS_PER_CTA consecutive ray groups; 2 CTAs are resident per SM and the
    // hardware scheduler balances the rest
    const long long blocks = (groups + RR_GROUPS_PER_CTA - 1) / RR_GROUPS_PER_CTA;
    if (blocks > 0x7fffffffLL) return RTXNERF_EINVAL;
    if (render)
        rays_roles_kernel<true><<<(unsigned)blocks, RR_THREADS, RrSmem::BYTES, stream>>>(
            ll, lv, reinterpret_cast<const uint32_t*>(enc->table), mlp->weights, a);
    else
        rays_roles_kernel<false><<<(unsigned)blocks, RR_THREADS, RrSmem::BYTES, stream>>>(
            ll, lv, reinterpret_cast<const uint32_t*>(enc->table), mlp->weights, a);
    return rtx_launch_status();
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
