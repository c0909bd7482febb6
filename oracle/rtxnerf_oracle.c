/*
 * rtxnerf_oracle.c — CPU restatement of the rtx_nerf per-ray hot path.
 * TEST INFRASTRUCTURE ONLY (see rtxnerf_oracle.h).  Compile with
 *   gcc -O3 -march=x86-64-v3 -ffp-contract=off -fopenmp -shared -fPIC
 * -ffp-contract=off is REQUIRED: every fused multiply-add the reference's PTX
 * contains is an explicit fmaf() below, and nothing else may fuse.
 */
#include "rtxnerf_oracle.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>

#ifdef _OPENMP
#include <omp.h>
#endif

/* ------------------------------------------------------------------------- */
/* PTX min.f32 / max.f32 semantics (what CUDA's min/max(float,float) compile to,
 * optixPrograms.cu:145-156,207): a NaN operand yields the other operand. */
static inline float cu_fminf(float a, float b)
{
    if (a != a) return b;
    if (b != b) return a;
    if (a == b) return signbit(a) ? a : b; /* -0 < +0 */
    return a < b ? a : b;
}
static inline float cu_fmaxf(float a, float b)
{
    if (a != a) return b;
    if (b != b) return a;
    if (a == b) return signbit(a) ? b : a;
    return a > b ? a : b;
}

/* ------------------------------------------------------------------------- */
/* fp16 <-> fp32, round to nearest even (cvt.rn.f16.f32 / cvt.f32.f16). */
uint16_t orc_f2h(float f)
{
    _Float16 h = (_Float16)f;
    uint16_t u;
    memcpy(&u, &h, 2);
    return u;
}
float orc_h2f(uint16_t u)
{
    _Float16 h;
    memcpy(&h, &u, 2);
    return (float)h;
}

/* ------------------------------------------------------------------------- */
/* main.cu:154-174 */
void orc_make_grid(int G, float* aabb6)
{
    float box_length = 2.0f / (float)G;
    long long p = 0;
    for (int x = 0; x < G; x++)
        for (int y = 0; y < G; y++)
            for (int z = 0; z < G; z++, p++) {
                float* a = aabb6 + 6 * p;
                a[0] = -1.0f + (float)x * box_length;
                a[3] = -1.0f + x * box_length + box_length;
                a[1] = -1.0f + y * box_length;
                a[4] = -1.0f + y * box_length + box_length;
                a[2] = -1.0f + z * box_length;
                a[5] = -1.0f + z * box_length + box_length;
            }
}

/* ------------------------------------------------------------------------- */
/* optixPrograms.cu:43-82.  Contraction pattern read from the reference's PTX:
 *   xd = fma(l0,u, l1*v) - l2*focal ;  norm = sqrt(fma(zd,zd, fma(xd,xd, yd*yd))) */
void orc_raygen(const float* l, float focal, float aspect, unsigned width, unsigned height,
                unsigned x, unsigned y, float o[3], float d[3], float vd[2])
{
    /* :56-57 — evaluated in double (the literals 0.5 and 2 promote), then rounded */
    double ud = (double)x + 0.5;
    ud = ud + ud;
    ud = ud / (double)width;
    ud = ud + -1.0;
    ud = ud * (double)aspect;
    float u = (float)ud;
    double vdbl = (double)y + 0.5;
    vdbl = vdbl + vdbl;
    vdbl = vdbl / (double)height;
    vdbl = vdbl + -1.0;
    float v = (float)vdbl;

    /* :62-64 */
    float xd = fmaf(l[0], u, l[1] * v) - l[2] * focal;
    float yd = fmaf(l[4], u, l[5] * v) - l[6] * focal;
    float zd = fmaf(l[8], u, l[9] * v) - l[10] * focal;
    /* :66-69 */
    float norm = sqrtf(fmaf(zd, zd, fmaf(xd, xd, yd * yd)));
    xd /= norm;
    yd /= norm;
    zd /= norm;
    /* :71-73 */
    vd[0] = atan2f(sqrtf(fmaf(xd, xd, yd * yd)), zd);
    vd[1] = atan2f(yd, xd);
    /* :75-78 */
    o[0] = l[3] / 10.0f;
    o[1] = l[7] / 10.0f;
    o[2] = l[11] / 10.0f;
    d[0] = xd;
    d[1] = yd;
    d[2] = zd;
}

/* optixPrograms.cu:79-80: tmin = 0, tmax = (max_point.z - min_point.z) + 100 = 102 */
#define RAY_TMIN 0.0f
#define RAY_TMAX 102.0f

/* optixPrograms.cu:132-169 on bounds (lo,hi) per axis.  Returns 1 and *t_rep when the
 * program calls optixReportIntersection. */
static inline int slab_report(const float lo[3], const float hi[3], const float o[3],
                              const float d[3], float* t_rep)
{
    float tmin = -INFINITY, tmax = INFINITY;
    for (int k = 0; k < 3; k++) {
        float t1 = (lo[k] - o[k]) / d[k];
        float t2 = (hi[k] - o[k]) / d[k];
        tmin = cu_fmaxf(tmin, cu_fminf(t1, t2));
        tmax = cu_fminf(tmax, cu_fmaxf(t1, t2));
    }
    if (tmax > tmin) {
        if (tmin < 0 && (double)tmax > 1e-6) tmin = 0;
        *t_rep = tmin;
        return 1;
    }
    return 0;
}

/* optixPrograms.cu:180-248 */
static inline void closest_hit(const float lo[3], const float hi[3], const float o[3],
                               const float d[3], float t_hit, float start[3], float end[3],
                               float* t_e_out)
{
    float tk[3];
    for (int k = 0; k < 3; k++) {
        start[k] = fmaf(t_hit, d[k], o[k]);           /* :194-196 */
        float plane = d[k] < 0 ? lo[k] : hi[k];        /* :201-203 */
        tk[k] = (plane - o[k]) / d[k];                 /* :204-206 */
    }
    float t_e = cu_fminf(cu_fminf(tk[0], tk[1]), tk[2]); /* :207 */
    for (int k = 0; k < 3; k++) end[k] = fmaf(d[k], t_e, o[k]); /* :209-211 */
    *t_e_out = t_e;
}

static inline void emit_hit(const orc_march_out* out, long long slot, float t_hit, float t_e,
                            const float start[3], const float end[3], int prim)
{
    if (out->t_start) out->t_start[slot] = t_hit;
    if (out->t_end) out->t_end[slot] = t_e;
    if (out->start_points) memcpy(out->start_points + 3 * slot, start, 12);
    if (out->end_points) memcpy(out->end_points + 3 * slot, end, 12);
    if (out->hit_prims) out->hit_prims[slot] = prim;
}

static inline void emit_ray(const orc_march_out* out, unsigned ray, const float o[3],
                            const float vd[2])
{
    if (out->view_dirs) {
        out->view_dirs[2 * (size_t)ray] = vd[0];
        out->view_dirs[2 * (size_t)ray + 1] = vd[1];
    }
    if (out->ray_origins) memcpy(out->ray_origins + 3 * (size_t)ray, o, 12);
}

static inline int out_cap(const orc_march_out* out)
{
    return out->max_hits > 0 ? out->max_hits : out->arr_size;
}

/* ------------------------------------------------------------------------- */
int orc_ray_march_brute(const float* look_at, float focal, float aspect, unsigned width,
                        unsigned height, const float* aabb6, const int* prim_ids, int n_prims,
                        unsigned ray_begin, unsigned ray_end, orc_march_out* out, int nthreads)
{
    int capped = 0;
    const int cap = out_cap(out);
    (void)height;
#pragma omp parallel for schedule(dynamic, 16) num_threads(nthreads > 0 ? nthreads : 1) reduction(+ : capped)
    for (long long r = ray_begin; r < (long long)ray_end; r++) {
        unsigned ray = (unsigned)r;
        unsigned x = ray % width, y = ray / width; /* :45 */
        float o[3], d[3], vd[2];
        orc_raygen(look_at, focal, aspect, width, height, x, y, o, d, vd);
        emit_ray(out, ray, o, vd);
        long long base = out->indices ? out->indices[ray] : (long long)ray * out->arr_size;
        int nh = 0;
        for (;;) { /* :99-115 relaunch loop */
            if (nh >= cap) {
                capped++;
                break;
            }
            float best_t = RAY_TMAX;
            int best = -1;
            for (int p = 0; p < n_prims; p++) {
                const float* b = aabb6 + 6 * (size_t)p;
                float t;
                if (!slab_report(b, b + 3, o, d, &t)) continue;
                /* OptiX accepts rayTmin <= t <= current rayTmax; equal-t ties keep the
                 * lowest primitive index (declared convention). */
                if (t >= RAY_TMIN && (best < 0 ? t <= best_t : t < best_t)) {
                    best_t = t;
                    best = p;
                }
            }
            if (best < 0) break; /* :171-178 miss */
            const float* b = aabb6 + 6 * (size_t)best;
            float start[3], end[3], t_e;
            closest_hit(b, b + 3, o, d, best_t, start, end, &t_e);
            emit_hit(out, base + nh, best_t, t_e, start, end, prim_ids ? prim_ids[best] : best);
            nh++;
            memcpy(o, end, 12); /* :245-247 */
        }
        out->num_hits[ray] = nh;
    }
    return capped;
}

/* ------------------------------------------------------------------------- */
/* Exact interval sweep over a power-of-two grid.
 *
 * For G = 2^k the make_grid planes p(i) = -1 + i*(2/G) are exact and shared between
 * neighbouring cells, so per axis the slab intervals [te_k(i), tx_k(i)] computed with
 * the reference's own expression (plane - o)/d tile the t line in index order.  The
 * cells whose three intervals overlap with positive length — the only ones the
 * reference predicate (tmax > tmin) can report — are visited, in increasing tmin, by
 * always advancing the axis whose interval ends first.  Every visited occupied cell is
 * then judged by the literal predicate, so the sweep only proposes; it never decides. */
static inline int locate_cell(float ok, int G, float L, float half_g)
{
    int i = (int)floorf((ok + 1.0f) * half_g);
    if (i < 0) i = 0;
    if (i > G - 1) i = G - 1;
    while (i > 0 && ok < -1.0f + (float)i * L) i--;
    while (i < G - 1 && ok >= -1.0f + (float)(i + 1) * L) i++;
    return i;
}

static int march_dda_ray(const float o0[3], const float d[3], const uint8_t* occ, int G,
                         int cap, const orc_march_out* out, long long base, int* capped)
{
    const float L = 2.0f / (float)G;
    const float half_g = 0.5f * (float)G;
    float o[3] = {o0[0], o0[1], o0[2]};
    int nh = 0;
    for (;;) {
        if (nh >= cap) {
            (*capped)++;
            break;
        }
        int i[3];
        float tl[3], th[3];
        for (int k = 0; k < 3; k++) {
            i[k] = locate_cell(o[k], G, L, half_g);
            tl[k] = ((-1.0f + (float)i[k] * L) - o[k]) / d[k];
            th[k] = ((-1.0f + (float)(i[k] + 1) * L) - o[k]) / d[k];
        }
        int found = 0;
        for (;;) {
            if ((unsigned)i[0] >= (unsigned)G || (unsigned)i[1] >= (unsigned)G ||
                (unsigned)i[2] >= (unsigned)G)
                break;
            float tx[3];
            float tmin = -INFINITY, tmax = INFINITY;
            for (int k = 0; k < 3; k++) {
                tmin = cu_fmaxf(tmin, cu_fminf(tl[k], th[k]));
                tx[k] = cu_fmaxf(tl[k], th[k]);
                tmax = cu_fminf(tmax, tx[k]);
            }
            long long p = ((long long)i[0] * G + i[1]) * G + i[2];
            if (occ[p] && tmax > tmin) {
                float t = tmin;
                if (tmin < 0 && (double)tmax > 1e-6) t = 0;
                if (t >= RAY_TMIN && t <= RAY_TMAX) {
                    float lo[3], hi[3], start[3], end[3], t_e;
                    for (int k = 0; k < 3; k++) {
                        lo[k] = -1.0f + (float)i[k] * L;
                        hi[k] = -1.0f + (float)(i[k] + 1) * L;
                    }
                    closest_hit(lo, hi, o, d, t, start, end, &t_e);
                    emit_hit(out, base + nh, t, t_e, start, end, (int)p);
                    nh++;
                    memcpy(o, end, 12);
                    found = 1;
                    break;
                }
            }
            /* advance the axis whose interval ends first (NaN / +inf never advance) */
            int ks = -1;
            float best = INFINITY;
            for (int k = 0; k < 3; k++)
                if (tx[k] < best) {
                    best = tx[k];
                    ks = k;
                }
            if (ks < 0) break;          /* nothing can advance */
            if (best > RAY_TMAX) break; /* every later cell starts beyond the ray */
            if (d[ks] > 0) {
                i[ks]++;
                tl[ks] = th[ks];
                th[ks] = ((-1.0f + (float)(i[ks] + 1) * L) - o[ks]) / d[ks];
            } else {
                i[ks]--;
                th[ks] = tl[ks];
                tl[ks] = ((-1.0f + (float)i[ks] * L) - o[ks]) / d[ks];
            }
        }
        if (!found) break;
    }
    return nh;
}

int orc_ray_march_dda(const float* look_at, float focal, float aspect, unsigned width,
                      unsigned height, const uint8_t* occ, int G, unsigned ray_begin,
                      unsigned ray_end, orc_march_out* out, int nthreads)
{
    if (G <= 0 || (G & (G - 1))) return -1;
    int capped = 0;
    const int cap = out_cap(out);
#pragma omp parallel for schedule(dynamic, 64) num_threads(nthreads > 0 ? nthreads : 1) reduction(+ : capped)
    for (long long r = ray_begin; r < (long long)ray_end; r++) {
        unsigned ray = (unsigned)r;
        unsigned x = ray % width, y = ray / width;
        float o[3], d[3], vd[2];
        orc_raygen(look_at, focal, aspect, width, height, x, y, o, d, vd);
        emit_ray(out, ray, o, vd);
        long long base = out->indices ? out->indices[ray] : (long long)ray * out->arr_size;
        int c = 0;
        out->num_hits[ray] = march_dda_ray(o, d, occ, G, cap, out, base, &c);
        capped += c;
    }
    return capped;
}

/* ------------------------------------------------------------------------- */
/* main.cu:631-637 */
long long orc_exclusive_scan(const int* num_hits, int n, int* indices)
{
    long long acc = 0;
    for (int i = 0; i < n; i++) {
        indices[i] = (int)acc;
        acc += num_hits[i];
    }
    return acc;
}

/* main.cu:646-676 */
void orc_compact_hits(const int* num_hits, const int* indices, int n_rays, int arr_size,
                      const float* dense, int comps, float* compact)
{
    for (int r = 0; r < n_rays; r++)
        memcpy(compact + (size_t)indices[r] * comps, dense + (size_t)r * arr_size * comps,
               sizeof(float) * comps * (size_t)num_hits[r]);
}

/* ------------------------------------------------------------------------- */
/* thrust::minstd_rand (linear_congruential_engine<uint32,48271,0,2147483647>, default
 * seed 1) + thrust::uniform_real_distribution<float> as used at sampler.cu:69-70,83-84.
 * The generator is a by-value kernel argument (sampler.cu:25,117), so every thread
 * (ray) starts from the same state. */
typedef struct {
    uint32_t x;
} minstd;
static inline uint32_t minstd_next(minstd* g)
{
    g->x = (uint32_t)(((uint64_t)g->x * 48271u) % 2147483647u);
    return g->x;
}
static inline float uniform_real(minstd* g, float a, float b)
{
    float result = (float)(minstd_next(g) - 1u);
    result /= (1.0f + (float)(2147483646u - 1u));
    return fmaf(result, b - a, a);
}

/* Counter-based uniform draw in [0,1) for the seeded sampling modes: splitmix64 of
 * (seed, global sample index), top 24 bits.  Unlike the reference's by-value minstd_rand
 * (sampler.cu:25,117: every thread draws the same sequence) each sample gets its own draw and a
 * launch is reproducible from (seed, index) alone. */
float orc_counter_uniform(unsigned long long seed, unsigned long long index)
{
    unsigned long long z = seed + (index + 1ull) * 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    z = z ^ (z >> 31);
    return (float)(z >> 40) * (1.0f / 16777216.0f);
}

void orc_sampler_seeded(const float* start_points, const float* end_points, const float* view_dirs,
                        float* t_vals, float* samples, int batch_size, const int* num_hits,
                        const int* indices, int S, int sample_type, unsigned long long seed, int nthreads);

/* sampler/sampler.cu:14-103 */
void orc_sampler(const float* start_points, const float* end_points, const float* view_dirs,
                 float* t_vals, float* samples, int batch_size, const int* num_hits,
                 const int* indices, int S, int sample_type, int nthreads)
{
    orc_sampler_seeded(start_points, end_points, view_dirs, t_vals, samples, batch_size, num_hits, indices, S,
                       sample_type, 0ull, nthreads);
}

/* seed == 0: the reference's own draws (by-value default-seeded minstd_rand); seed != 0: the
 * same modes with counter-based draws, t = fma(u, b - a, a). */
void orc_sampler_seeded(const float* start_points, const float* end_points, const float* view_dirs,
                        float* t_vals, float* samples, int batch_size, const int* num_hits,
                        const int* indices, int S, int sample_type, unsigned long long seed, int nthreads)
{
    const float inv_s = 1.0f / (float)S; /* sampler.cu:50,65 */
#pragma omp parallel for schedule(dynamic, 256) num_threads(nthreads > 0 ? nthreads : 1)
    for (int x = 0; x < batch_size; x++) {
        int start_index = indices[x];
        int n_hits = num_hits[x];
        float theta = view_dirs[2 * (size_t)x], phi = view_dirs[2 * (size_t)x + 1];
        minstd rng = {1u};
        for (int j = 0; j < n_hits; j++) {
            const float* origin = start_points + 3 * (size_t)(start_index + j);
            const float* finish = end_points + 3 * (size_t)(start_index + j);
            float dir[3] = {finish[0] - origin[0], finish[1] - origin[1], finish[2] - origin[2]};
            float t_initial = 0.0f, t_final = inv_s;
            for (int i = 0; i < S; i++) {
                size_t n = (size_t)(start_index + j) * S + i;
                float* s = samples + 5 * n;
                float t;
                if (sample_type == 0) { /* :54-66 */
                    t = t_initial;
                    t_initial += inv_s;
                    t_vals[n] = t_initial;
                } else if (sample_type == 2) { /* :68-81 */
                    t = seed ? orc_counter_uniform(seed, n) : uniform_real(&rng, 0.0f, 1.0f);
                    t_vals[n] = t_initial;
                } else { /* :82-99 */
                    t = seed ? fmaf(orc_counter_uniform(seed, n), t_final - t_initial, t_initial)
                             : uniform_real(&rng, t_initial, t_final);
                    t_vals[n] = t_initial;
                    t_initial = t_final;
                    t_final += inv_s;
                }
                s[0] = fmaf(t, dir[0], origin[0]);
                s[1] = fmaf(t, dir[1], origin[1]);
                s[2] = fmaf(t, dir[2], origin[2]);
                s[3] = theta;
                s[4] = phi;
            }
        }
    }
}

/* ------------------------------------------------------------------------- */
void orc_cell_radiance(const int* hit_prims, long long n_hits, int S, const float* cell_rgbs4,
                       float* radiance)
{
    for (long long h = 0; h < n_hits; h++) {
        const float* c = cell_rgbs4 + 4 * (size_t)hit_prims[h];
        for (int i = 0; i < S; i++) memcpy(radiance + 4 * ((size_t)h * S + i), c, 16);
    }
}

/* ------------------------------------------------------------------------- */
/* vol_render/vol_render.cu:19-73.  PTX contraction: T = fma(sigma,delta,T);
 * acc = fma(c, e^-T * (1 - e^(sigma*-delta)), acc). */
void orc_volrender(const float* network_outputs, const int* num_hits, const int* indices,
                   const float* ray_hit, int batch_size, int S, float* pixels, int nthreads)
{
#pragma omp parallel for schedule(dynamic, 256) num_threads(nthreads > 0 ? nthreads : 1)
    for (int ray = 0; ray < batch_size; ray++) {
        int start_index = indices[ray];
        int n = num_hits[ray];
        float transmittance = 0.0f, t_initial = 0.0f;
        float acc[3] = {0.0f, 0.0f, 0.0f};
        for (int j = 0; j < n; j++)
            for (int i = 0; i < S; i++) {
                size_t s = (size_t)(start_index + j) * S + i;
                const float* c = network_outputs + 4 * s; /* :50-53 */
                float sigma = c[3];
                float t = ray_hit[s];                     /* :55 */
                float delta = fabsf(t - t_initial);       /* :56 */
                t_initial = t;
                transmittance = fmaf(sigma, delta, transmittance); /* :60 */
                float w = expf(-transmittance) * (1.0f - expf(sigma * -delta)); /* :61-63 */
                acc[0] = fmaf(c[0], w, acc[0]);           /* :65-67 */
                acc[1] = fmaf(c[1], w, acc[1]);
                acc[2] = fmaf(c[2], w, acc[2]);
            }
        pixels[3 * (size_t)ray] = acc[0];
        pixels[3 * (size_t)ray + 1] = acc[1];
        pixels[3 * (size_t)ray + 2] = acc[2];
    }
}

/* vol_render/vol_render.cu:75-143 (not the analytic gradient; reproduced literally). */
void orc_volrender_backward(const uint16_t* loss_gradients, const float* radiance,
                            const float* t_hit, const int* num_hits, const int* indices,
                            int batch_size, int S, uint16_t* radiance_gradients, int nthreads)
{
#pragma omp parallel for schedule(dynamic, 256) num_threads(nthreads > 0 ? nthreads : 1)
    for (int ray = 0; ray < batch_size; ray++) {
        int start_index = indices[ray];
        int n = num_hits[ray];
        float t_initial = 0.0f;
        float lg[3] = {orc_h2f(loss_gradients[3 * (size_t)ray]),
                       orc_h2f(loss_gradients[3 * (size_t)ray + 1]),
                       orc_h2f(loss_gradients[3 * (size_t)ray + 2])}; /* :121-123 */
        for (int j = 0; j < n; j++)
            for (int i = 0; i < S; i++) {
                size_t s = (size_t)(start_index + j) * S + i;
                const float* c = radiance + 4 * s;
                float sigma = c[3];
                float t = t_hit[s];
                float delta = fabsf(t - t_initial); /* :115 */
                t_initial = t;
                float tp = sigma * delta;           /* :118 */
                float a[3] = {lg[0] * tp, tp * lg[1], tp * lg[2]};
                float e = expf(delta * -sigma);     /* exp(-sigma*delta) :127 */
                float ds = fmaf(delta * (c[0] * a[0]), e, 0.0f); /* :127-129 */
                ds = fmaf(delta * (c[1] * a[1]), e, ds);
                ds = fmaf(delta * (c[2] * a[2]), e, ds);
                float om = 1.0f - expf(sigma * -delta); /* :133-135 */
                uint16_t* g = radiance_gradients + 4 * s;
                g[0] = orc_f2h(a[0] * om);
                g[1] = orc_f2h(a[1] * om);
                g[2] = orc_f2h(a[2] * om);
                g[3] = orc_f2h(ds);
            }
    }
}

/* ------------------------------------------------------------------------- */
/* Multiresolution hash encoding — restated from the published instant-ngp / tiny-cuda-nn
 * algorithm (tcnn @ d0522b4 is absent: PARITY UNPINNED).  Call site: main.cu:721. */
static inline void level_geometry(const orc_hashgrid_cfg* cfg, int level, float* scale,
                                  uint32_t* res, uint32_t* size)
{
    /* scale = b^level * N_min - 1 ; res = ceil(scale) + 1 */
    double s = exp2((double)level * log2((double)cfg->per_level_scale)) * (double)cfg->base_resolution - 1.0;
    *scale = (float)s;
    *res = (uint32_t)ceilf(*scale) + 1u;
    uint64_t dense = (uint64_t)(*res) * (*res) * (*res);
    dense = (dense + 7u) / 8u * 8u;
    uint64_t cap = 1ull << cfg->log2_hashmap_size;
    *size = (uint32_t)(dense < cap ? dense : cap);
}

long long orc_hashgrid_offsets(const orc_hashgrid_cfg* cfg, long long* offsets)
{
    long long acc = 0;
    for (int l = 0; l < cfg->n_levels; l++) {
        float scale;
        uint32_t res, size;
        level_geometry(cfg, l, &scale, &res, &size);
        offsets[l] = acc;
        acc += size;
    }
    offsets[cfg->n_levels] = acc;
    return acc;
}

void orc_hashgrid_geometry(const orc_hashgrid_cfg* cfg, float* scales, uint32_t* ress, uint32_t* sizes)
{
    for (int l = 0; l < cfg->n_levels; l++) level_geometry(cfg, l, &scales[l], &ress[l], &sizes[l]);
}

static inline uint32_t grid_index(uint32_t x, uint32_t y, uint32_t z, uint32_t res, uint32_t size)
{
    uint64_t dense = (uint64_t)res * res * res;
    uint32_t idx;
    if (dense <= size)
        idx = x + y * res + z * res * res;
    else
        idx = (x * 1u) ^ (y * 2654435761u) ^ (z * 805459861u);
    return idx % size;
}

/* fp16 fused multiply-add with one rounding, for operands that are fp16 values held in floats
 * (HFMA2's per-lane semantics).  a*b is exact in fp32 (two 11-bit significands); the sum with c
 * is rounded to fp32 TO ODD (TwoSum gives the exact error, the sticky bit goes into the last
 * place), and a value rounded to odd at 24 bits rounds to fp16's 11 bits exactly as the
 * infinitely precise one does.  Same results as (_Float16)((double)a*b + c) at a tenth of the
 * cost: double -> fp16 is a library call on x86, fp32 -> fp16 is one F16C instruction. */
static inline _Float16 hfma_rn(float a, float b, float c)
{
    const float p = a * b; /* exact */
    const float s = p + c;
    const float bv = s - p;
    const float err = (p - (s - bv)) + (c - bv); /* TwoSum: p + c = s + err exactly */
    uint32_t bits;
    memcpy(&bits, &s, 4);
    if (err != 0.0f && !(bits & 1u) && s == s && (bits & 0x7f800000u) != 0x7f800000u) {
        const int same_sign = (err > 0.0f) == (s > 0.0f || (s == 0.0f && err > 0.0f));
        if (s == 0.0f) {
            /* |true value| = |err| > 0 below the smallest fp32 subnormal step of s: not reachable for fp16 inputs */
        } else if (same_sign) {
            bits |= 1u;
        } else {
            bits -= 1u;
        }
    }
    float so;
    memcpy(&so, &bits, 4);
    return (_Float16)so;
}

/* [-1,1] -> [0,1] with the clamp of fma.rn.sat.f32 (hashgrid.cuh unit_coord): a position outside
 * the cube takes the features of its nearest face, a NaN those of 0, and no corner index can leave
 * its level (tiny-cuda-nn keeps indices in range with its modulo; this is the cheaper guard). */
static inline float unit_coord(float p)
{
    float x = fmaf(p, 0.5f, 0.5f);
    if (!(x >= 0.0f)) x = 0.0f;
    if (x > 1.0f) x = 1.0f;
    return x;
}

void orc_hashgrid_encode(int L, int F, const float* scales, const uint32_t* ress,
                         const uint32_t* sizes, const long long* offsets, const uint16_t* table,
                         const float* samples, int sample_stride, long long n, uint16_t* encoded,
                         int nthreads)
{
#pragma omp parallel for schedule(static) num_threads(nthreads > 0 ? nthreads : 1)
    for (long long s = 0; s < n; s++) {
        const float* p = samples + (size_t)s * sample_stride;
        float x01[3];
        for (int k = 0; k < 3; k++) x01[k] = unit_coord(p[k]); /* [-1,1] -> [0,1], clamped */
        for (int l = 0; l < L; l++) {
            float frac[3];
            uint32_t g[3];
            for (int k = 0; k < 3; k++) {
                float pos = fmaf(scales[l], x01[k], 0.5f);
                float fl = floorf(pos);
                frac[k] = pos - fl;
                g[k] = (uint32_t)(int)fl;
            }
            /* tiny-cuda-nn's N-linear interpolation (grid kernel): weight in fp32, cast to the
             * parameter type, result = fma((T)weight, grid_val, result) in the parameter type
             * (T = __half), corners in order c = cx | cy<<1 | cz<<2.  fma in fp16 = ONE rounding
             * of the exact w*v + acc (hfma_rn above). */
            _Float16 acc[8] = {0};
            for (int c = 0; c < 8; c++) {
                float w = 1.0f;
                uint32_t q[3];
                for (int k = 0; k < 3; k++) {
                    if (c & (1 << k)) {
                        w *= frac[k];
                        q[k] = g[k] + 1u;
                    } else {
                        w *= 1.0f - frac[k];
                        q[k] = g[k];
                    }
                }
                uint32_t idx = grid_index(q[0], q[1], q[2], ress[l], sizes[l]);
                const uint16_t* e = table + ((size_t)offsets[l] + idx) * F;
                const float wh = (float)(_Float16)w;
                for (int f = 0; f < F; f++) acc[f] = hfma_rn(wh, orc_h2f(e[f]), (float)acc[f]);
            }
            for (int f = 0; f < F; f++) {
                uint16_t bits;
                memcpy(&bits, &acc[f], 2);
                encoded[(size_t)s * L * F + (size_t)l * F + f] = bits;
            }
        }
    }
}

/* FullyFusedMLP forward — restated from the published description (bias-free, fp16
 * weights and activations, fp32 accumulate, output padded to 16).  PARITY UNPINNED.
 * Call site: main.cu:721; output slice + half->float: main.cu:723-728. */
/* ---------------------------------------------------------------------------------------
 * Frequency (positional) encoding of the reference's literal network config, main.cu:47-61:
 * Composite[Frequency(3 dims, n_frequencies 10) + Frequency(2 dims; the "n_bins" key is not a
 * Frequency parameter, so tiny-cuda-nn's default of 12 frequencies applies)], output padded
 * to a multiple of 16 with ones (108 -> 112).  tiny-cuda-nn is absent from the archive
 * (PARITY UNPINNED); the published definition is restated: for input dim d and k < F
 *   out[d*2F + 2k] = sin(pi * 2^k * x_d),  out[d*2F + 2k + 1] = cos(pi * 2^k * x_d)
 * evaluated in double here and rounded once to fp16.  The kernel reduces the argument
 * exactly (mod 2) and uses the hardware sine: equal up to one fp16 ulp.
 * ------------------------------------------------------------------------------------- */
void orc_frequency_encode(const float* samples, int stride, long long n, int n_freq_xyz, int n_freq_dir,
                          int padded_width, uint16_t* out, int nthreads)
{
    const double PI = 3.14159265358979323846;
#pragma omp parallel for schedule(static) num_threads(nthreads > 0 ? nthreads : 1)
    for (long long s = 0; s < n; s++) {
        const float* p = samples + (size_t)s * stride;
        uint16_t* o = out + (size_t)s * padded_width;
        int j = 0;
        for (int d = 0; d < 5; d++) {
            const int F = d < 3 ? n_freq_xyz : n_freq_dir;
            for (int k = 0; k < F; k++) {
                const double y = ldexp((double)p[d], k);
                o[j++] = orc_f2h((float)sin(PI * y));
                o[j++] = orc_f2h((float)cos(PI * y));
            }
        }
        for (; j < padded_width; j++) o[j] = orc_f2h(1.0f);
    }
}

void orc_mlp_forward(const uint16_t* weights, int in_dim, int width, int n_hidden,
                     int out_activation, const uint16_t* in, long long n, uint16_t* out16,
                     float* radiance4, int nthreads)
{
    /* layer shapes: [width][in_dim], (n_hidden-1) x [width][width], [16][width] */
    size_t wsz = (size_t)width * in_dim + (size_t)(n_hidden - 1) * width * width + (size_t)16 * width;
    float* wf = (float*)malloc(sizeof(float) * wsz);
    for (size_t i = 0; i < wsz; i++) wf[i] = orc_h2f(weights[i]);
#pragma omp parallel num_threads(nthreads > 0 ? nthreads : 1)
    {
        float* a = (float*)malloc(sizeof(float) * (size_t)(width > in_dim ? width : in_dim));
        float* b = (float*)malloc(sizeof(float) * (size_t)width);
#pragma omp for schedule(static)
        for (long long s = 0; s < n; s++) {
            for (int k = 0; k < in_dim; k++) a[k] = orc_h2f(in[(size_t)s * in_dim + k]);
            const float* w = wf;
            int cur = in_dim;
            for (int layer = 0; layer < n_hidden; layer++) {
                for (int r = 0; r < width; r++) {
                    double acc = 0.0;
                    for (int k = 0; k < cur; k++) acc += (double)w[(size_t)r * cur + k] * (double)a[k];
                    float v = (float)acc;
                    v = v > 0.0f ? v : 0.0f;
                    b[r] = orc_h2f(orc_f2h(v));
                }
                memcpy(a, b, sizeof(float) * (size_t)width);
                w += (size_t)width * cur;
                cur = width;
            }
            for (int r = 0; r < 16; r++) {
                double acc = 0.0;
                for (int k = 0; k < cur; k++) acc += (double)w[(size_t)r * cur + k] * (double)a[k];
                float v = (float)acc;
                if (out_activation == 1) v = 1.0f / (1.0f + expf(-v));
                uint16_t h = orc_f2h(v);
                if (out16) out16[(size_t)s * 16 + r] = h;
                if (radiance4 && r < 4) radiance4[(size_t)s * 4 + r] = orc_h2f(h);
            }
        }
        free(a);
        free(b);
    }
    free(wf);
}

/* ---------------------------------------------------------------------------------------
 * Training step restatements (tiny-cuda-nn network->backward main.cu:781, loss->evaluate
 * main.cu:755-760, optimizer->step main.cu:787 — tcnn absent: PARITY UNPINNED).  They follow
 * the rounding points of the CUDA path (fp16 activations / activation gradients, wide
 * accumulation) so that the comparison tolerance only has to absorb summation order.
 * ------------------------------------------------------------------------------------- */
void orc_mlp_backward(const uint16_t* weights, int out_activation, const uint16_t* enc,
                      const uint16_t* dl_dradiance, long long n, uint16_t* d_enc, double* grad_w)
{
    enum { IN = 32, W = 64, OUT = 16 };
    const size_t o1 = (size_t)W * IN, o2 = o1 + (size_t)W * W, wsz = o2 + (size_t)OUT * W;
    float* wf = (float*)malloc(sizeof(float) * wsz);
    for (size_t i = 0; i < wsz; i++) wf[i] = orc_h2f(weights[i]);
    const float *w0 = wf, *w1 = wf + o1, *w2 = wf + o2;
    memset(grad_w, 0, sizeof(double) * wsz);
    for (long long s = 0; s < n; s++) {
        float a0[IN], h1[W], h2[W], dout[OUT], dh2[W], dh1[W];
        for (int k = 0; k < IN; k++) a0[k] = orc_h2f(enc[(size_t)s * IN + k]);
        for (int r = 0; r < W; r++) {
            double acc = 0.0;
            for (int k = 0; k < IN; k++) acc += (double)w0[r * IN + k] * a0[k];
            float v = (float)acc;
            h1[r] = orc_h2f(orc_f2h(v > 0.f ? v : 0.f));
        }
        for (int r = 0; r < W; r++) {
            double acc = 0.0;
            for (int k = 0; k < W; k++) acc += (double)w1[r * W + k] * h1[k];
            float v = (float)acc;
            h2[r] = orc_h2f(orc_f2h(v > 0.f ? v : 0.f));
        }
        for (int r = 0; r < OUT; r++) {
            dout[r] = 0.f;
            if (r >= 4) continue;
            double acc = 0.0;
            for (int k = 0; k < W; k++) acc += (double)w2[r * W + k] * h2[k];
            float da = 1.0f;
            if (out_activation == 1) {
                float sg = 1.0f / (1.0f + expf(-(float)acc));
                da = sg * (1.0f - sg);
            }
            dout[r] = orc_h2f(orc_f2h(orc_h2f(dl_dradiance[(size_t)s * 4 + r]) * da));
        }
        for (int r = 0; r < OUT; r++)
            for (int k = 0; k < W; k++) grad_w[o2 + (size_t)r * W + k] += (double)dout[r] * h2[k];
        for (int k = 0; k < W; k++) {
            double acc = 0.0;
            for (int r = 0; r < OUT; r++) acc += (double)dout[r] * w2[r * W + k];
            dh2[k] = orc_h2f(orc_f2h(h2[k] > 0.f ? (float)acc : 0.f));
        }
        for (int r = 0; r < W; r++)
            for (int k = 0; k < W; k++) grad_w[o1 + (size_t)r * W + k] += (double)dh2[r] * h1[k];
        for (int k = 0; k < W; k++) {
            double acc = 0.0;
            for (int r = 0; r < W; r++) acc += (double)dh2[r] * w1[r * W + k];
            dh1[k] = orc_h2f(orc_f2h(h1[k] > 0.f ? (float)acc : 0.f));
        }
        for (int r = 0; r < W; r++)
            for (int k = 0; k < IN; k++) grad_w[(size_t)r * IN + k] += (double)dh1[r] * a0[k];
        for (int k = 0; k < IN; k++) {
            double acc = 0.0;
            for (int r = 0; r < W; r++) acc += (double)dh1[r] * w0[r * IN + k];
            d_enc[(size_t)s * IN + k] = orc_f2h((float)acc);
        }
    }
    free(wf);
}

/* Backward of the reference's literal network (Composite Frequency -> 8 x 128 ReLU -> 16, main.cu:47-68,781),
 * weight gradients only (the encoding has no parameters).  Same rounding points as the CUDA path:
 * fp16 activations, fp16 dOut, fp16 activation gradients after the ReLU mask, wide accumulation.
 * weights: row-major [128][112], 7 x [128][128], [16][128]; enc: [n][112] fp16 (orc_frequency_encode).
 * PARITY UNPINNED (tiny-cuda-nn absent); pinned to fp64 autograd in tests/test_oracle_autograd.py. */
void orc_freqnet_backward(const uint16_t* weights, int out_activation, const uint16_t* enc,
                          const uint16_t* dl_dradiance, long long n, double* grad_w)
{
    enum { IN = 112, W = 128, H = 8, OUT = 16 };
    const size_t wsz = (size_t)W * IN + (size_t)(H - 1) * W * W + (size_t)OUT * W;
    float* wf = (float*)malloc(sizeof(float) * wsz);
    for (size_t i = 0; i < wsz; i++) wf[i] = orc_h2f(weights[i]);
    size_t off[H + 1];
    off[0] = 0;
    off[1] = (size_t)W * IN;
    for (int l = 2; l <= H; l++) off[l] = off[l - 1] + (size_t)W * W;
    memset(grad_w, 0, sizeof(double) * wsz);
    float (*x)[W] = malloc(sizeof(float) * W * (H + 1));
    for (long long s = 0; s < n; s++) {
        /* forward: x[0] = encoding (112), x[l+1] = fp16(relu(W_l x[l])) */
        for (int k = 0; k < IN; k++) x[0][k] = orc_h2f(enc[(size_t)s * IN + k]);
        for (int l = 0; l < H; l++) {
            const int K = l == 0 ? IN : W;
            const float* w = wf + off[l];
            for (int r = 0; r < W; r++) {
                double acc = 0.0;
                for (int k = 0; k < K; k++) acc += (double)w[(size_t)r * K + k] * x[l][k];
                float v = (float)acc;
                x[l + 1][r] = orc_h2f(orc_f2h(v > 0.f ? v : 0.f));
            }
        }
        float dz[W], dprev[W], dout[OUT];
        const float* w8 = wf + off[H];
        for (int r = 0; r < OUT; r++) {
            dout[r] = 0.f;
            if (r >= 4) continue;
            double acc = 0.0;
            for (int k = 0; k < W; k++) acc += (double)w8[(size_t)r * W + k] * x[H][k];
            float da = 1.0f;
            if (out_activation == 1) {
                float sg = 1.0f / (1.0f + expf(-(float)acc));
                da = sg * (1.0f - sg);
            }
            dout[r] = orc_h2f(orc_f2h(orc_h2f(dl_dradiance[(size_t)s * 4 + r]) * da));
        }
        for (int r = 0; r < OUT; r++)
            for (int k = 0; k < W; k++) grad_w[off[H] + (size_t)r * W + k] += (double)dout[r] * x[H][k];
        for (int k = 0; k < W; k++) {
            double acc = 0.0;
            for (int r = 0; r < OUT; r++) acc += (double)dout[r] * w8[(size_t)r * W + k];
            dz[k] = orc_h2f(orc_f2h(x[H][k] > 0.f ? (float)acc : 0.f));      /* dZ_7 */
        }
        for (int l = H - 1; l >= 0; l--) {
            const int K = l == 0 ? IN : W;
            const float* w = wf + off[l];
            for (int r = 0; r < W; r++)
                for (int k = 0; k < K; k++) grad_w[off[l] + (size_t)r * K + k] += (double)dz[r] * x[l][k];
            if (l == 0) break;
            for (int k = 0; k < W; k++) {
                double acc = 0.0;
                for (int r = 0; r < W; r++) acc += (double)dz[r] * w[(size_t)r * K + k];
                dprev[k] = orc_h2f(orc_f2h(x[l][k] > 0.f ? (float)acc : 0.f));
            }
            memcpy(dz, dprev, sizeof(dz));
        }
    }
    free(x);
    free(wf);
}

/* dL/d table: every sample adds w_corner * dEnc[level] to its 8 corner entries (F = 2). */
void orc_hashgrid_backward(int L, const float* scales, const uint32_t* ress, const uint32_t* sizes,
                           const long long* offsets, const float* samples, int sample_stride,
                           long long n, const uint16_t* d_enc, double* grad_table)
{
    for (long long s = 0; s < n; s++) {
        const float* p = samples + (size_t)s * sample_stride;
        float x01[3];
        for (int k = 0; k < 3; k++) x01[k] = unit_coord(p[k]); /* [-1,1] -> [0,1], clamped */
        for (int l = 0; l < L; l++) {
            float g0 = orc_h2f(d_enc[(size_t)s * L * 2 + (size_t)l * 2]);
            float g1 = orc_h2f(d_enc[(size_t)s * L * 2 + (size_t)l * 2 + 1]);
            float frac[3];
            uint32_t g[3];
            for (int k = 0; k < 3; k++) {
                float pos = fmaf(scales[l], x01[k], 0.5f);
                float fl = floorf(pos);
                frac[k] = pos - fl;
                g[k] = (uint32_t)(int)fl;
            }
            for (int c = 0; c < 8; c++) {
                float w = 1.0f;
                uint32_t q[3];
                for (int k = 0; k < 3; k++) {
                    if (c & (1 << k)) { w *= frac[k]; q[k] = g[k] + 1u; }
                    else { w *= 1.0f - frac[k]; q[k] = g[k]; }
                }
                uint32_t idx = grid_index(q[0], q[1], q[2], ress[l], sizes[l]);
                double* e = grad_table + ((size_t)offsets[l] + idx) * 2;
                e[0] += (double)(w * g0);
                e[1] += (double)(w * g1);
            }
        }
    }
}
