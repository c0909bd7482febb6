// march.cu — occupancy-grid ray march for sm_100a.
//
// Replaces the reference's OptiX pipeline (rtx/src/optixPrograms.cu:43-248, launched at
// main.cu:506-508) — B200 has no RT cores.  Semantics reproduced bit for bit:
//   raygen           optixPrograms.cu:43-82   (u,v in double; d normalised with IEEE sqrt/div)
//   relaunch loop    optixPrograms.cu:99-115  (next origin = previous exit point, t local to it)
//   slab predicate   optixPrograms.cu:132-169 (tmax > tmin; clamp to 0 when tmin<0 && tmax>1e-6)
//   acceptance       ray interval [0, 102]    (optixPrograms.cu:79-80)
//   closest hit      optixPrograms.cu:180-248 (entry/exit points, t_start/t_end, num_hits++)
//
// Method: for a power-of-two G the make_grid planes (main.cu:154-174) are exact and shared
// by neighbouring cells, so per axis the slab intervals computed with the reference's own
// expression (plane - o)/d tile the t line.  The cells the reference predicate can report
// are exactly those whose three intervals overlap with positive length; they are visited
// in increasing tmin by always advancing the axis whose interval ends first (a 3-way
// sorted-interval sweep = an exact 3D-DDA).  Every visited OCCUPIED cell is then judged by
// the literal predicate: the sweep proposes, the reference's arithmetic decides.
// All arithmetic that feeds a comparison uses the _rn intrinsics (common.cuh).
#include "common.cuh"

#include <math.h>
#include <stdlib.h>
#include <new>
#include <vector>

#define RAY_TMAX 102.0f              // optixPrograms.cu:80: (max.z - min.z) + 100
#define CLAMP_EPS 9.99999997e-07f    // (float)1e-6: for float t, (double)t > 1e-6  <=>  t > (float)1e-6
#define MARCH_BLOCK 128

// ------------------------------------------------------------------------------------ raygen
struct RayGen {
    float ox, oy, oz, dx, dy, dz, theta, phi;
};

// optixPrograms.cu:43-82; FMA placement follows the reference's PTX:
//   xd = fma(l0,u, l1*v) - l2*focal ; norm = sqrt(fma(zd,zd, fma(xd,xd, yd*yd)))
__device__ __forceinline__ RayGen raygen(const float* __restrict__ l, float focal, float aspect,
                                         unsigned width, unsigned height, unsigned x, unsigned y)
{
    double ud = __dadd_rn((double)x, 0.5);
    ud = __dadd_rn(ud, ud);
    ud = __ddiv_rn(ud, (double)width);
    ud = __dadd_rn(ud, -1.0);
    ud = __dmul_rn(ud, (double)aspect);
    float u = __double2float_rn(ud);
    double vd = __dadd_rn((double)y, 0.5);
    vd = __dadd_rn(vd, vd);
    vd = __ddiv_rn(vd, (double)height);
    vd = __dadd_rn(vd, -1.0);
    float v = __double2float_rn(vd);

    float xd = rn_sub(rn_fma(l[0], u, rn_mul(l[1], v)), rn_mul(l[2], focal));
    float yd = rn_sub(rn_fma(l[4], u, rn_mul(l[5], v)), rn_mul(l[6], focal));
    float zd = rn_sub(rn_fma(l[8], u, rn_mul(l[9], v)), rn_mul(l[10], focal));
    float norm = __fsqrt_rn(rn_fma(zd, zd, rn_fma(xd, xd, rn_mul(yd, yd))));
    RayGen r;
    r.dx = rn_div(xd, norm);
    r.dy = rn_div(yd, norm);
    r.dz = rn_div(zd, norm);
    r.theta = atan2f(__fsqrt_rn(rn_fma(r.dx, r.dx, rn_mul(r.dy, r.dy))), r.dz);
    r.phi = atan2f(r.dy, r.dx);
    r.ox = rn_div(l[3], 10.0f);
    r.oy = rn_div(l[7], 10.0f);
    r.oz = rn_div(l[11], 10.0f);
    return r;
}

// ------------------------------------------------------------------------------------ occupancy bits
// Two copies of the same G^3 bits live in a grid: the linear x-major array of the ABI (bit p & 31 of word p >> 5,
// p = (x*G+y)*G+z) and a BRICK-ordered one, a 64-bit word per 4x4x4 cells (bit (x&3)<<4 | (y&3)<<2 | z&3 of
// word ((x>>2)*GB + (y>>2))*GB + (z>>2), GB = ceil(G/4)), rebuilt on the device whenever the bits change.  The
// sweep keeps the brick it is in in registers: one load serves every step the ray spends inside those 64 cells
// (2.5-4 consecutive steps whatever the occupancy) and a 128-byte line covers a 3-D block of cells instead of
// 1024 cells along z, where the linear order costs one L2-latency load per visited cell.  Which cells are
// visited and how each is judged is unchanged — the hit lists are the same bits.
struct GridBits {
    const uint32_t* lin;
    const unsigned long long* bricks;   // null: linear bits only (RTXNERF_MARCH_BRICKS=0)
    unsigned gb;
};
struct BrickCache {
    unsigned id;                        // brick whose word is cached; ~0u: none
    unsigned long long word;
};
__device__ __forceinline__ bool cell_occupied(const GridBits& b, int G, unsigned x, unsigned y, unsigned z, BrickCache& c)
{
    if (b.bricks) {
        const unsigned id = ((x >> 2) * b.gb + (y >> 2)) * b.gb + (z >> 2);
        if (id != c.id) {
            c.id = id;
            c.word = __ldg(b.bricks + id);
        }
        return (c.word >> (((x & 3u) << 4) | ((y & 3u) << 2) | (z & 3u))) & 1ull;
    }
    const unsigned p = (x * (unsigned)G + y) * (unsigned)G + z;
    return (__ldg(b.lin + (p >> 5)) >> (p & 31u)) & 1u;
}

__global__ void brick_build_kernel(const uint32_t* __restrict__ lin, int G, unsigned gb, unsigned long long* __restrict__ bricks)
{
    const long long id = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (id >= (long long)gb * gb * gb) return;
    const unsigned bz = (unsigned)(id % gb), by = (unsigned)((id / gb) % gb), bx = (unsigned)(id / ((long long)gb * gb));
    unsigned long long w = 0ull;
    for (unsigned k = 0; k < 64u; k++) {
        const unsigned x = 4u * bx + (k >> 4), y = 4u * by + ((k >> 2) & 3u), z = 4u * bz + (k & 3u);
        if (x < (unsigned)G && y < (unsigned)G && z < (unsigned)G) {
            const unsigned long long p = ((unsigned long long)x * G + y) * G + z;
            w |= (unsigned long long)((__ldg(lin + (p >> 5)) >> (p & 31u)) & 1u) << k;
        }
    }
    bricks[id] = w;
}

// ------------------------------------------------------------------------------------ sweep
__device__ __forceinline__ float plane_at(int i, float L) { return rn_fma((float)i, L, -1.0f); }  // exact

struct HitRecord {
    float t_start, t_end;
    float sx, sy, sz, ex, ey, ez;
    int cell;
};

// IEEE-754 a/d (round to nearest even) with the reciprocal hoisted out of the march loop.
// A ray divides by the same three direction components hundreds of times
// (optixPrograms.cu:140-151,204-206 evaluate (plane - o)/d for every candidate cell);
// div.rn.f32 costs MUFU.RCP + 2 refinement FMAs + FCHK + branch + 5 FMAs each time.  With
// r = RN(1/d) computed once per ray, two Markstein correction steps give the correctly
// rounded quotient whenever no intermediate can underflow or overflow: the residual
// a - d*q is an exact multiple of 2^(exp(a)-46), representable for |a| >= 2^-103.  FAST
// certifies the operand ranges (per ray for d, per relaunch for o: dir_fast/origin_fast);
// anything else takes div.rn itself.  a is never -0 here (a = plane - o with plane != -0),
// so the sign of a zero quotient is the IEEE one as well.  Checked against a/d on the host
// for 3e8 random and patterned operand pairs: 0 mismatches.
__device__ __forceinline__ float div_exact(float a, float d, float r)
{
    const float q0 = rn_mul(a, r);
    const float e0 = rn_fma(-d, q0, a);
    const float q1 = rn_fma(e0, r, q0);
    const float e1 = rn_fma(-d, q1, a);
    return rn_fma(e1, r, q1);
}

// d certifies the fast path when 2^-30 <= |d| < 4 (direction components are <= 1)
__device__ __forceinline__ bool dir_fast(float d)
{
    return ((__float_as_uint(d) >> 23) & 0xffu) - (127u - 30u) < 32u;
}
// o certifies it when o == 0 or 2^-90 <= |o| < 2^60: then a = plane - o (planes are
// multiples of 2/G, |plane| <= 1) is 0 or has 2^-103 <= |a| < 2^61
__device__ __forceinline__ bool origin_fast(float o)
{
    const uint32_t b = __float_as_uint(o);
    return (((b >> 23) & 0xffu) - (127u - 90u) < 150u) | ((b << 1) == 0u);
}

// Per-axis sweep state.  near/far are the two slab values of the current cell labelled by the
// sign of d (far = (lo-o)/d when d<0 else (hi-o)/d, the value optixPrograms.cu:201-206 picks as
// the exit candidate); {near, far} == {tl, th}, so min/max of the pair are the reference's.
struct AxisState {
    float near_t, far_t;   // slab values of the current cell
    float far_plane;       // plane the far value belongs to (exact float)
    int i;                 // cell index on this axis
};

__device__ __forceinline__ void axis_enter(AxisState& a, float o, float d, float r, int G, float L, float half_g)
{
    // cell containing o (planes compared exactly), clamped to the grid.  The estimate e = floor(RN(o+1) * G/2)
    // is never below the true index and at most one above it: o >= plane_i <=> o+1 >= i*L exactly, RN is monotone
    // and i*L is representable, so RN(o+1) >= i*L; and o < plane_(i+1) gives RN(o+1) <= (i+1)*L, equality only by
    // rounding up.  One conditional decrement therefore lands on the cell locate_cell's loops find (outside the
    // grid the clamp decides; certified origins are finite).
    int i = (int)floorf(rn_mul(rn_add(o, 1.0f), half_g));
    i = max(0, min(G - 1, i));
    if (i > 0 && o < plane_at(i, L)) i--;
    const float lo = plane_at(i, L), hi = plane_at(i + 1, L);
    const float tl = div_exact(rn_sub(lo, o), d, r), th = div_exact(rn_sub(hi, o), d, r);
    const bool neg = d < 0.0f;
    a.i = i;
    a.near_t = neg ? th : tl;
    a.far_t = neg ? tl : th;
    a.far_plane = neg ? lo : hi;
}

// One launch of the reference's relaunch loop (optixPrograms.cu:99-115): sweep from origin o
// until a cell is accepted (emit + true) or the ray leaves the grid / interval (false).
// Certified operands only (every d normal and non-zero, every slab value finite): the sweep
// direction and the exit candidate both follow the sign of d, so the state is (near, far).
template <bool WANT_START, class Emit>
__device__ __forceinline__ bool march_segment(const GridBits& bits, int G, float L, float half_g,
                                              float& ox, float& oy, float& oz, float dx, float dy, float dz,
                                              float rx, float ry, float rz, int& nh, Emit& emit)
{
    AxisState X, Y, Z;
    axis_enter(X, ox, dx, rx, G, L, half_g);
    axis_enter(Y, oy, dy, ry, G, L, half_g);
    axis_enter(Z, oz, dz, rz, G, L, half_g);
    // signed plane step per axis; the sign of d is fixed for the ray (d == 0 never advances)
    const float slx = dx < 0.0f ? -L : L, sly = dy < 0.0f ? -L : L, slz = dz < 0.0f ? -L : L;
    const int six = dx < 0.0f ? -1 : 1, siy = dy < 0.0f ? -1 : 1, siz = dz < 0.0f ? -1 : 1;
    BrickCache brick{~0u, 0ull};
    for (;;) {
        // optixPrograms.cu:143-156 (fminf/fmaxf == PTX min/max: a NaN operand yields the other)
        const float txx = fmaxf(X.near_t, X.far_t), txy = fmaxf(Y.near_t, Y.far_t), txz = fmaxf(Z.near_t, Z.far_t);
        const bool occ = cell_occupied(bits, G, (unsigned)X.i, (unsigned)Y.i, (unsigned)Z.i, brick);
        if (occ) {
            const unsigned p = ((unsigned)X.i * (unsigned)G + (unsigned)Y.i) * (unsigned)G + (unsigned)Z.i;
            const float tmin = fmaxf(fmaxf(fmaxf(-INFINITY, fminf(X.near_t, X.far_t)), fminf(Y.near_t, Y.far_t)),
                                     fminf(Z.near_t, Z.far_t));
            const float tmax = fminf(fminf(fminf(INFINITY, txx), txy), txz);
            if (tmax > tmin) {                                              // :164
                const float t = (tmin < 0.0f && tmax > CLAMP_EPS) ? 0.0f : tmin;  // :165-166
                if (t >= 0.0f && t <= RAY_TMAX) {                          // OptiX ray interval
                    HitRecord h;
                    h.t_start = t;                                          // :188-189
                    if (WANT_START) {
                        h.sx = rn_fma(t, dx, ox);                           // :194-196
                        h.sy = rn_fma(t, dy, oy);
                        h.sz = rn_fma(t, dz, oz);
                    }
                    // :201-207 exit plane by the sign of d = the far value of each axis
                    const float t_e = fminf(fminf(X.far_t, Y.far_t), Z.far_t);
                    h.t_end = t_e;
                    h.ex = rn_fma(dx, t_e, ox);                             // :209-211
                    h.ey = rn_fma(dy, t_e, oy);
                    h.ez = rn_fma(dz, t_e, oz);
                    h.cell = (int)p;
                    emit(nh, h);
                    nh++;                                                   // :241
                    ox = h.ex;                                              // :245-247 relaunch from the exit point
                    oy = h.ey;
                    oz = h.ez;
                    return true;
                }
            }
        }
        // advance the axis whose interval ends first (ties: lowest axis; NaN/inf never advance)
        int ks = -1;
        float best = INFINITY;
        if (txx < best) { best = txx; ks = 0; }
        if (txy < best) { best = txy; ks = 1; }
        if (txz < best) { best = txz; ks = 2; }
        if (ks < 0 || best > RAY_TMAX) return false;
        const float os = ks == 0 ? ox : (ks == 1 ? oy : oz);
        const float ds = ks == 0 ? dx : (ks == 1 ? dy : dz);
        const float rs = ks == 0 ? rx : (ks == 1 ? ry : rz);
        const float pl = (ks == 0 ? X.far_plane : (ks == 1 ? Y.far_plane : Z.far_plane)) +
                         (ks == 0 ? slx : (ks == 1 ? sly : slz));          // next plane: exact
        const float tn = div_exact(rn_sub(pl, os), ds, rs);
        int moved;
        if (ks == 0) { X.i += six; X.near_t = X.far_t; X.far_t = tn; X.far_plane = pl; moved = X.i; }
        else if (ks == 1) { Y.i += siy; Y.near_t = Y.far_t; Y.far_t = tn; Y.far_plane = pl; moved = Y.i; }
        else { Z.i += siz; Z.near_t = Z.far_t; Z.far_t = tn; Z.far_plane = pl; moved = Z.i; }
        if ((unsigned)moved >= (unsigned)G) return false;                   // left the grid
    }
}

// The same launch for operands outside the certified range (zero / denormal / huge origin
// components, |d| < 2^-30, d == +-0): div.rn everywhere, (tl, th) state, and the sweep
// direction taken from `ds > 0` exactly as the oracle does, so the degenerate cases
// (inf / NaN slab values) resolve identically.
__device__ __forceinline__ int locate_cell(float ok, int G, float L, float half_g)
{
    int i = (int)floorf(rn_mul(rn_add(ok, 1.0f), half_g));
    i = max(0, min(G - 1, i));
    while (i > 0 && ok < plane_at(i, L)) i--;
    while (i < G - 1 && ok >= plane_at(i + 1, L)) i++;
    return i;
}

template <bool WANT_START, class Emit>
__device__ __noinline__ bool march_segment_generic(const GridBits& bits, int G, float L, float half_g,
                                                   float& ox, float& oy, float& oz, float dx, float dy, float dz,
                                                   int& nh, Emit& emit)
{
    int ix = locate_cell(ox, G, L, half_g);
    int iy = locate_cell(oy, G, L, half_g);
    int iz = locate_cell(oz, G, L, half_g);
    float tlx = rn_div(rn_sub(plane_at(ix, L), ox), dx), thx = rn_div(rn_sub(plane_at(ix + 1, L), ox), dx);
    float tly = rn_div(rn_sub(plane_at(iy, L), oy), dy), thy = rn_div(rn_sub(plane_at(iy + 1, L), oy), dy);
    float tlz = rn_div(rn_sub(plane_at(iz, L), oz), dz), thz = rn_div(rn_sub(plane_at(iz + 1, L), oz), dz);
    BrickCache brick{~0u, 0ull};
    for (;;) {
        if ((unsigned)ix >= (unsigned)G || (unsigned)iy >= (unsigned)G || (unsigned)iz >= (unsigned)G) return false;
        float txx = fmaxf(tlx, thx), txy = fmaxf(tly, thy), txz = fmaxf(tlz, thz);
        float tmin = fmaxf(fmaxf(fmaxf(-INFINITY, fminf(tlx, thx)), fminf(tly, thy)), fminf(tlz, thz));
        float tmax = fminf(fminf(fminf(INFINITY, txx), txy), txz);
        unsigned p = ((unsigned)ix * (unsigned)G + (unsigned)iy) * (unsigned)G + (unsigned)iz;
        bool occ = cell_occupied(bits, G, (unsigned)ix, (unsigned)iy, (unsigned)iz, brick);
        if (occ && tmax > tmin) {
            float t = (tmin < 0.0f && tmax > CLAMP_EPS) ? 0.0f : tmin;
            if (t >= 0.0f && t <= RAY_TMAX) {
                HitRecord h;
                h.t_start = t;
                if (WANT_START) {
                    h.sx = rn_fma(t, dx, ox);
                    h.sy = rn_fma(t, dy, oy);
                    h.sz = rn_fma(t, dz, oz);
                }
                float t_e = fminf(fminf(dx < 0.0f ? tlx : thx, dy < 0.0f ? tly : thy), dz < 0.0f ? tlz : thz);
                h.t_end = t_e;
                h.ex = rn_fma(dx, t_e, ox);
                h.ey = rn_fma(dy, t_e, oy);
                h.ez = rn_fma(dz, t_e, oz);
                h.cell = (int)p;
                emit(nh, h);
                nh++;
                ox = h.ex;
                oy = h.ey;
                oz = h.ez;
                return true;
            }
        }
        int ks = -1;
        float best = INFINITY;
        if (txx < best) { best = txx; ks = 0; }
        if (txy < best) { best = txy; ks = 1; }
        if (txz < best) { best = txz; ks = 2; }
        if (ks < 0 || best > RAY_TMAX) return false;
        float os = ks == 0 ? ox : (ks == 1 ? oy : oz);
        float ds = ks == 0 ? dx : (ks == 1 ? dy : dz);
        int is = ks == 0 ? ix : (ks == 1 ? iy : iz);
        float tls = ks == 0 ? tlx : (ks == 1 ? tly : tlz);
        float ths = ks == 0 ? thx : (ks == 1 ? thy : thz);
        if (ds > 0.0f) {
            is++;
            tls = ths;
            ths = rn_div(rn_sub(plane_at(is + 1, L), os), ds);
        } else {
            is--;
            ths = tls;
            tls = rn_div(rn_sub(plane_at(is, L), os), ds);
        }
        if (ks == 0) { ix = is; tlx = tls; thx = ths; }
        else if (ks == 1) { iy = is; tly = tls; thy = ths; }
        else { iz = is; tlz = tls; thz = ths; }
    }
}

// Marches one ray; calls emit(j, HitRecord) for hit j.  Returns the number of hits (<= cap).
template <bool WANT_START, class Emit>
__device__ __forceinline__ int march_ray(const GridBits& bits, int G, float ox, float oy,
                                         float oz, float dx, float dy, float dz, int cap, Emit emit)
{
    const float L = 2.0f / (float)G;
    const float half_g = 0.5f * (float)G;
    const float rx = __frcp_rn(dx), ry = __frcp_rn(dy), rz = __frcp_rn(dz);
    const bool dfast = dir_fast(dx) & dir_fast(dy) & dir_fast(dz);
    int nh = 0;
    while (nh < cap) {
        const bool fast = dfast & origin_fast(ox) & origin_fast(oy) & origin_fast(oz);
        bool found;
        if (fast)
            found = march_segment<WANT_START>(bits, G, L, half_g, ox, oy, oz, dx, dy, dz, rx, ry, rz, nh, emit);
        else
            found = march_segment_generic<WANT_START>(bits, G, L, half_g, ox, oy, oz, dx, dy, dz, nh, emit);
        if (!found) break;
    }
    return nh;
}

// The same march as a FLAT loop: one sweep step per iteration for every lane that is inside a
// segment, and the relaunch work (three exact cell locates + six exact divisions, ~150
// instructions) batched: lanes that have just emitted a hit wait until the relaunch block runs
// for the warp, which happens when every live lane wants it, when at least ENTER_MIN lanes do,
// or every ENTER_EVERY iterations.  march_ray lets each lane run to its next hit and the warp
// then waits for the slowest lane (with scattered occupancy ~38 steps for a mean of ~10); here
// a lane idles for at most ENTER_EVERY - 1 steps.  The operations of a ray and their order are
// unchanged, so the records are the same bits.  All lanes of `mask` must call it together.
// Measured (round h): config 5 (512^3 Bernoulli 0.10, H_max 64) staged march 18.8 -> 13.5 ms;
// config 2 (lego, 2.3 steps per hit, H_max 16) 0.33 -> 0.41 ms, so the shared-memory one-pass
// path keeps march_ray and the staged path (max_hits > 32) takes this one.
__constant__ int c_enter_every = 4, c_enter_min = 12;

template <bool WANT_START, class Emit>
__device__ __forceinline__ int march_ray_flat(const GridBits& bits, int G, float ox, float oy,
                                              float oz, float dx, float dy, float dz, int cap, bool valid,
                                              unsigned mask, Emit emit)
{
    const float L = 2.0f / (float)G;
    const float half_g = 0.5f * (float)G;
    const float rx = __frcp_rn(dx), ry = __frcp_rn(dy), rz = __frcp_rn(dz);
    const bool dfast = dir_fast(dx) & dir_fast(dy) & dir_fast(dz);
    const float slx = dx < 0.0f ? -L : L, sly = dy < 0.0f ? -L : L, slz = dz < 0.0f ? -L : L;
    const int six = dx < 0.0f ? -1 : 1, siy = dy < 0.0f ? -1 : 1, siz = dz < 0.0f ? -1 : 1;
    const int enter_every = c_enter_every, enter_min = c_enter_min;
    AxisState X, Y, Z;
    X.near_t = X.far_t = X.far_plane = 0.f; X.i = 0;
    Y = X; Z = X;
    int nh = 0, since = 0;
    BrickCache brick{~0u, 0ull};        // kept across the segments of the ray
    bool done = !valid || cap <= 0, need_enter = true;
    for (;;) {
        const unsigned live = __ballot_sync(mask, !done);
        if (!live) break;
        const unsigned want = __ballot_sync(mask, need_enter && !done);
        if (want && (want == live || __popc(want) >= enter_min || since >= enter_every)) {
            since = 0;
            if (need_enter && !done) {
                if (dfast & origin_fast(ox) & origin_fast(oy) & origin_fast(oz)) {
                    axis_enter(X, ox, dx, rx, G, L, half_g);
                    axis_enter(Y, oy, dy, ry, G, L, half_g);
                    axis_enter(Z, oz, dz, rz, G, L, half_g);
                    need_enter = false;
                } else {      // uncertified operands: the whole segment through the generic code
                    const bool found = march_segment_generic<WANT_START>(bits, G, L, half_g, ox, oy, oz, dx, dy, dz,
                                                                         nh, emit);
                    if (!found || nh >= cap) done = true;
                }
            }
        } else {
            since++;
        }
        if (!need_enter && !done) {
            // one step of march_segment
            const float txx = fmaxf(X.near_t, X.far_t), txy = fmaxf(Y.near_t, Y.far_t), txz = fmaxf(Z.near_t, Z.far_t);
            const bool occ = cell_occupied(bits, G, (unsigned)X.i, (unsigned)Y.i, (unsigned)Z.i, brick);
            bool accepted = false;
            if (occ) {
                const unsigned p = ((unsigned)X.i * (unsigned)G + (unsigned)Y.i) * (unsigned)G + (unsigned)Z.i;
                const float tmin = fmaxf(fmaxf(fmaxf(-INFINITY, fminf(X.near_t, X.far_t)), fminf(Y.near_t, Y.far_t)),
                                         fminf(Z.near_t, Z.far_t));
                const float tmax = fminf(fminf(fminf(INFINITY, txx), txy), txz);
                if (tmax > tmin) {
                    const float t = (tmin < 0.0f && tmax > CLAMP_EPS) ? 0.0f : tmin;
                    if (t >= 0.0f && t <= RAY_TMAX) {
                        HitRecord h;
                        h.t_start = t;
                        if (WANT_START) {
                            h.sx = rn_fma(t, dx, ox);
                            h.sy = rn_fma(t, dy, oy);
                            h.sz = rn_fma(t, dz, oz);
                        }
                        const float t_e = fminf(fminf(X.far_t, Y.far_t), Z.far_t);
                        h.t_end = t_e;
                        h.ex = rn_fma(dx, t_e, ox);
                        h.ey = rn_fma(dy, t_e, oy);
                        h.ez = rn_fma(dz, t_e, oz);
                        h.cell = (int)p;
                        emit(nh, h);
                        nh++;
                        ox = h.ex;
                        oy = h.ey;
                        oz = h.ez;
                        accepted = true;
                        need_enter = true;
                        if (nh >= cap) done = true;
                    }
                }
            }
            if (!accepted) {
                int ks = -1;
                float best = INFINITY;
                if (txx < best) { best = txx; ks = 0; }
                if (txy < best) { best = txy; ks = 1; }
                if (txz < best) { best = txz; ks = 2; }
                if (ks < 0 || best > RAY_TMAX) {
                    done = true;
                } else {
                    const float os = ks == 0 ? ox : (ks == 1 ? oy : oz);
                    const float ds = ks == 0 ? dx : (ks == 1 ? dy : dz);
                    const float rs = ks == 0 ? rx : (ks == 1 ? ry : rz);
                    const float pl = (ks == 0 ? X.far_plane : (ks == 1 ? Y.far_plane : Z.far_plane)) +
                                     (ks == 0 ? slx : (ks == 1 ? sly : slz));
                    const float tn = div_exact(rn_sub(pl, os), ds, rs);
                    int moved;
                    if (ks == 0) { X.i += six; X.near_t = X.far_t; X.far_t = tn; X.far_plane = pl; moved = X.i; }
                    else if (ks == 1) { Y.i += siy; Y.near_t = Y.far_t; Y.far_t = tn; Y.far_plane = pl; moved = Y.i; }
                    else { Z.i += siz; Z.near_t = Z.far_t; Z.far_t = tn; Z.far_plane = pl; moved = Z.i; }
                    if ((unsigned)moved >= (unsigned)G) done = true;
                }
            }
        }
    }
    return nh;
}

// ------------------------------------------------------------------------------------ kernels
struct CamArgs {
    const float* look_at;
    float focal, aspect;
    unsigned width, height;
};

// Pass 1: count hits per ray, write per-ray outputs and one partial sum per block.
__global__ void __launch_bounds__(MARCH_BLOCK)
march_count_kernel(GridBits bits, int G, CamArgs cam, long long ray_begin, long long n_rays,
                   int cap, int* __restrict__ num_hits, float* __restrict__ view_dirs,
                   float* __restrict__ ray_origins, int* __restrict__ block_sums)
{
    __shared__ int warp_sums[MARCH_BLOCK / 32];
    long long lr = (long long)blockIdx.x * MARCH_BLOCK + threadIdx.x;
    int nh = 0;
    if (lr < n_rays) {
        unsigned long long ray = (unsigned long long)(ray_begin + lr);
        RayGen r = raygen(cam.look_at, cam.focal, cam.aspect, cam.width, cam.height,
                          (unsigned)(ray % cam.width), (unsigned)(ray / cam.width));
        nh = march_ray<false>(bits, G, r.ox, r.oy, r.oz, r.dx, r.dy, r.dz, cap, [](int, const HitRecord&) {});
        num_hits[lr] = nh;
        reinterpret_cast<float2*>(view_dirs)[lr] = make_float2(r.theta, r.phi);
        if (ray_origins) {
            ray_origins[3 * lr] = r.ox;
            ray_origins[3 * lr + 1] = r.oy;
            ray_origins[3 * lr + 2] = r.oz;
        }
    }
    int s = nh;
#pragma unroll
    for (int o = 16; o; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if ((threadIdx.x & 31) == 0) warp_sums[threadIdx.x >> 5] = s;
    __syncthreads();
    if (threadIdx.x == 0) {
        int t = 0;
#pragma unroll
        for (int w = 0; w < MARCH_BLOCK / 32; w++) t += warp_sums[w];
        block_sums[blockIdx.x] = t;
    }
}

// Pass 2: exclusive scan of the block sums (one CTA; nb is a few thousand).
__global__ void __launch_bounds__(1024)
block_scan_kernel(const int* __restrict__ block_sums, int nb, long long* __restrict__ block_offsets,
                  long long* __restrict__ total)
{
    __shared__ long long warp_tot[32];
    __shared__ long long carry_s;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) carry_s = 0;
    __syncthreads();
    for (int base = 0; base < nb; base += 1024) {
        int i = base + threadIdx.x;
        long long v = i < nb ? block_sums[i] : 0;
        long long inc = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            long long n = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += n;
        }
        if (lane == 31) warp_tot[warp] = inc;
        __syncthreads();
        if (warp == 0) {
            long long w = warp_tot[lane];
            long long winc = w;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                long long n = __shfl_up_sync(0xffffffffu, winc, o);
                if (lane >= o) winc += n;
            }
            warp_tot[lane] = winc - w;  // exclusive
        }
        __syncthreads();
        long long carry = carry_s;
        if (i < nb) block_offsets[i] = carry + warp_tot[warp] + inc - v;
        __syncthreads();
        if (threadIdx.x == 1023) carry_s = carry + warp_tot[31] + inc;
        __syncthreads();
    }
    if (threadIdx.x == 0) *total = carry_s;
}

struct FillOut {
    float* t_start;
    float* t_end;
    float* start_points;
    float* end_points;
    int* hit_cells;
    long long capacity;
    float* seg8;      // optional [capacity][8]: (sx, sy, sz, ex | ey, ez, 0, 0), one 32-byte sector per hit
};

__device__ __forceinline__ void store_seg8(float* seg8, long long slot, float sx, float sy, float sz, float ex, float ey,
                                           float ez)
{
    float4* p = reinterpret_cast<float4*>(seg8) + 2 * slot;
    p[0] = make_float4(sx, sy, sz, ex);
    p[1] = make_float4(ey, ez, 0.0f, 0.0f);
}

__device__ __forceinline__ void store_hit(const FillOut& o, long long slot, const HitRecord& h)
{
    if (slot >= o.capacity) return;
    if (o.t_start) o.t_start[slot] = h.t_start;
    if (o.t_end) o.t_end[slot] = h.t_end;
    if (o.start_points) {      // (null with the packed records alone: rtxnerf_ray_march_csr_seg)
        float* s = o.start_points + 3 * slot;
        s[0] = h.sx; s[1] = h.sy; s[2] = h.sz;
        float* e = o.end_points + 3 * slot;
        e[0] = h.ex; e[1] = h.ey; e[2] = h.ez;
    }
    if (o.hit_cells) o.hit_cells[slot] = h.cell;
    if (o.seg8) store_seg8(o.seg8, slot, h.sx, h.sy, h.sz, h.ex, h.ey, h.ez);
}

// Pass 3: block-local scan of the counts + block offset = CSR index; march again and store.
__global__ void __launch_bounds__(MARCH_BLOCK)
march_fill_kernel(GridBits bits, int G, CamArgs cam, long long ray_begin, long long n_rays,
                  int cap, const int* __restrict__ num_hits, const long long* __restrict__ block_offsets,
                  int* __restrict__ indices, FillOut out)
{
    __shared__ int warp_sums[MARCH_BLOCK / 32];
    long long lr = (long long)blockIdx.x * MARCH_BLOCK + threadIdx.x;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int nh = lr < n_rays ? num_hits[lr] : 0;
    int inc = warp_incl_scan_i32(nh, lane);
    if (lane == 31) warp_sums[warp] = inc;
    __syncthreads();
    int woff = 0;
#pragma unroll
    for (int w = 0; w < MARCH_BLOCK / 32; w++) woff += (w < warp) ? warp_sums[w] : 0;
    long long base = block_offsets[blockIdx.x] + woff + inc - nh;
    if (lr >= n_rays) return;
    indices[lr] = (int)base;
    if (nh == 0) return;
    unsigned long long ray = (unsigned long long)(ray_begin + lr);
    RayGen r = raygen(cam.look_at, cam.focal, cam.aspect, cam.width, cam.height, (unsigned)(ray % cam.width),
                      (unsigned)(ray / cam.width));
    march_ray<true>(bits, G, r.ox, r.oy, r.oz, r.dx, r.dy, r.dz, nh,
                    [&](int j, const HitRecord& h) { store_hit(out, base + j, h); });
}

// ------------------------------------------------------------------------------------ one pass
// Single-pass CSR march for max_hits <= ONEPASS_MAX_CAP: every ray is marched ONCE.
//   1. a CTA takes the next tile of MARCH_BLOCK consecutive rays from a ticket counter
//      (tickets = launch order, so a tile only ever waits on tiles that are already running);
//   2. each thread marches its ray and parks 6 words per hit in shared memory
//      (t_start, t_end, exit point, cell; row stride 6*cap+1 words: conflict-free);
//   3. block scan of the counts, then a decoupled look-back over the per-tile status words
//      ([flag:2 | hit count:62], flag 1 = tile aggregate, 2 = inclusive prefix) gives the
//      tile's CSR base without a second kernel or a second march;
//   4. the tile's hits are contiguous in every CSR array, so the CTA copies them out with
//      slot = thread: fully coalesced stores (the two-pass fill wrote 9 scalars per hit at
//      64..192-byte strides between lanes: 593 MB of DRAM traffic for 368 MB of records,
//      profiles/r01_d_view_full.txt).  Entry points are rebuilt on the way out from the
//      previous exit point (start = fma(t_start, d, o_prev), optixPrograms.cu:194-196).
#define ONEPASS_MAX_CAP 32
#ifndef MARCH_STAGED_MIN_CTAS
#define MARCH_STAGED_MIN_CTAS 8      // the staged grid is sized for 8 CTAs per SM: hold the registers to 64
#endif
#define LOOKBACK_SPIN_LIMIT (1u << 24)

__device__ __forceinline__ unsigned long long ld_status(const unsigned long long* p)
{
    unsigned long long v;
    asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_status(unsigned long long* p, unsigned long long v)
{
    asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

// STAGED = false: hits parked in shared memory, one tile per CTA (cap <= ONEPASS_MAX_CAP).
// STAGED = true (ONEPASS_MAX_CAP < cap <= STAGED_MAX_CAP): the same kernel with the hit records
// parked in a per-CTA slab of the workspace in HBM/L2 and a persistent grid that keeps taking
// tiles, so the slabs stay a few hundred MB whatever the ray count; still ONE traversal per ray
// (count -> scan -> fill walks every ray twice).
template <bool STAGED, bool FLAT>
__global__ void __launch_bounds__(MARCH_BLOCK, STAGED ? MARCH_STAGED_MIN_CTAS : 1)
march_onepass_kernel(GridBits bits, int G, CamArgs cam, long long ray_begin, long long n_rays,
                     int cap, int* __restrict__ num_hits, float* __restrict__ view_dirs,
                     float* __restrict__ ray_origins, int* __restrict__ indices, FillOut out,
                     unsigned long long* __restrict__ status, unsigned long long* __restrict__ ticket,
                     long long* __restrict__ total_hits, int n_tiles, uint32_t* __restrict__ gstage)
{
    extern __shared__ __align__(16) uint32_t dyn[];
    const int row = STAGED ? 6 * cap : 6 * cap + 1;
    uint32_t* rec = STAGED ? gstage + (size_t)blockIdx.x * MARCH_BLOCK * row : dyn;   // [MARCH_BLOCK][row]
    float* pts = reinterpret_cast<float*>(dyn + (STAGED ? 0 : MARCH_BLOCK * row));    // [2][MARCH_BLOCK*3]
    float* dirs = pts + 2 * MARCH_BLOCK * 3;                    // [MARCH_BLOCK][3]
    int* excl = reinterpret_cast<int*>(dirs + MARCH_BLOCK * 3); // [MARCH_BLOCK + 1]
    uint8_t* owner = reinterpret_cast<uint8_t*>(excl + MARCH_BLOCK + 4);   // [MARCH_BLOCK * cap]: slot -> ray
    __shared__ int warp_sums[MARCH_BLOCK / 32];
    __shared__ int tile_s;
    __shared__ long long base_s;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

    __shared__ float origin_s[3];
  for (;;) {                                                    // STAGED: until the tickets run out
    if (tid == 0) tile_s = (int)atomicAdd(ticket, 1ull);
    __syncthreads();
    const int tile = tile_s;
    if (tile >= n_tiles) return;
    const long long lr = (long long)tile * MARCH_BLOCK + tid;

    // ---- 1+2: march, park the hits
    int nh = 0;
    float ox0 = 0.f, oy0 = 0.f, oz0 = 0.f;
    if (FLAT) {
        const bool valid = lr < n_rays;
        RayGen r;
        r.ox = r.oy = r.oz = 0.f;
        r.dx = r.dy = r.dz = 0.5f;
        r.theta = r.phi = 0.f;
        if (valid) {
            unsigned long long ray = (unsigned long long)(ray_begin + lr);
            r = raygen(cam.look_at, cam.focal, cam.aspect, cam.width, cam.height, (unsigned)(ray % cam.width),
                       (unsigned)(ray / cam.width));
            ox0 = r.ox; oy0 = r.oy; oz0 = r.oz;
            dirs[3 * tid] = r.dx; dirs[3 * tid + 1] = r.dy; dirs[3 * tid + 2] = r.dz;
        }
        uint32_t* my = rec + tid * row;
        nh = march_ray_flat<false>(bits, G, r.ox, r.oy, r.oz, r.dx, r.dy, r.dz, cap, valid, 0xffffffffu,
                                   [&](int j, const HitRecord& h) {
            uint32_t* w = my + 6 * j;
            w[0] = __float_as_uint(h.t_start);
            w[1] = __float_as_uint(h.t_end);
            w[2] = __float_as_uint(h.ex);
            w[3] = __float_as_uint(h.ey);
            w[4] = __float_as_uint(h.ez);
            w[5] = (uint32_t)h.cell;
        });
        if (valid) {
            num_hits[lr] = nh;
            reinterpret_cast<float2*>(view_dirs)[lr] = make_float2(r.theta, r.phi);
            if (ray_origins) {
                ray_origins[3 * lr] = r.ox;
                ray_origins[3 * lr + 1] = r.oy;
                ray_origins[3 * lr + 2] = r.oz;
            }
        }
    } else if (lr < n_rays) {
        unsigned long long ray = (unsigned long long)(ray_begin + lr);
        RayGen r = raygen(cam.look_at, cam.focal, cam.aspect, cam.width, cam.height,
                          (unsigned)(ray % cam.width), (unsigned)(ray / cam.width));
        ox0 = r.ox; oy0 = r.oy; oz0 = r.oz;
        dirs[3 * tid] = r.dx; dirs[3 * tid + 1] = r.dy; dirs[3 * tid + 2] = r.dz;
        uint32_t* my = rec + tid * row;
        nh = march_ray<false>(bits, G, r.ox, r.oy, r.oz, r.dx, r.dy, r.dz, cap, [&](int j, const HitRecord& h) {
            uint32_t* w = my + 6 * j;
            w[0] = __float_as_uint(h.t_start);
            w[1] = __float_as_uint(h.t_end);
            w[2] = __float_as_uint(h.ex);
            w[3] = __float_as_uint(h.ey);
            w[4] = __float_as_uint(h.ez);
            w[5] = (uint32_t)h.cell;
        });
        num_hits[lr] = nh;
        reinterpret_cast<float2*>(view_dirs)[lr] = make_float2(r.theta, r.phi);
        if (ray_origins) {
            ray_origins[3 * lr] = r.ox;
            ray_origins[3 * lr + 1] = r.oy;
            ray_origins[3 * lr + 2] = r.oz;
        }
    }
    // ---- 3: block scan + look-back
    const int inc = warp_incl_scan_i32(nh, lane);
    if (lane == 31) warp_sums[warp] = inc;
    __syncthreads();
    int woff = 0, block_total = 0;
#pragma unroll
    for (int w = 0; w < MARCH_BLOCK / 32; w++) {
        woff += (w < warp) ? warp_sums[w] : 0;
        block_total += warp_sums[w];
    }
    const int my_excl = woff + inc - nh;
    excl[tid] = my_excl;
    if (tid == 0) excl[MARCH_BLOCK] = block_total;
    for (int j = 0; j < nh; j++) owner[my_excl + j] = (uint8_t)tid;
    if (warp == 0) {
        long long exclusive = 0;
        if (tile > 0) {
            if (lane == 0) st_status(status + tile, (1ull << 62) | (unsigned long long)block_total);
            int idx = tile - 1;
            unsigned spins = 0;
            for (;;) {
                const int mine = idx - lane;
                unsigned long long v = 2ull << 62;                       // before tile 0: prefix 0
                if (mine >= 0) v = ld_status(status + mine);
                while (__any_sync(0xffffffffu, (v >> 62) == 0ull)) {
                    if (++spins > LOOKBACK_SPIN_LIMIT) __trap();         // never hang the GPU
                    __nanosleep(20);
                    if (mine >= 0 && (v >> 62) == 0ull) v = ld_status(status + mine);
                }
                const unsigned prefix_mask = __ballot_sync(0xffffffffu, (v >> 62) == 2ull);
                const int first = prefix_mask ? (__ffs(prefix_mask) - 1) : 31;
                long long contrib = (lane <= first) ? (long long)(v & ((1ull << 62) - 1ull)) : 0ll;
#pragma unroll
                for (int o = 16; o; o >>= 1) contrib += __shfl_xor_sync(0xffffffffu, contrib, o);
                exclusive += contrib;
                if (prefix_mask) break;
                idx -= 32;
            }
        }
        if (lane == 0) {
            st_status(status + tile, (2ull << 62) | (unsigned long long)(exclusive + block_total));
            base_s = exclusive;
            if (tile == n_tiles - 1) *total_hits = exclusive + block_total;
        }
    }
    __syncthreads();
    const long long base = base_s;
    if (lr < n_rays) indices[lr] = (int)(base + excl[tid]);
    // the origin is the same for every ray of a view (optixPrograms.cu:75-78); thread 0 of a
    // tile always holds a ray
    if (tid == 0) { origin_s[0] = ox0; origin_s[1] = oy0; origin_s[2] = oz0; }
    __syncthreads();
    const float o0x = origin_s[0], o0y = origin_s[1], o0z = origin_s[2];

    // ---- 4: copy out, slot = thread
    for (int k0 = 0; k0 < block_total; k0 += MARCH_BLOCK) {
        const int k = k0 + tid;
        const bool live = k < block_total;
        if (live) {
            const int lo = owner[k];                             // the ray that owns slot k
            const int j = k - excl[lo];
            const uint32_t* w = rec + lo * row + 6 * j;
            const float ts = __uint_as_float(w[0]), te = __uint_as_float(w[1]);
            const float ex = __uint_as_float(w[2]), ey = __uint_as_float(w[3]), ez = __uint_as_float(w[4]);
            float px = o0x, py = o0y, pz = o0z;                  // the origin this hit was launched from
            if (j > 0) { px = __uint_as_float(w[-4]); py = __uint_as_float(w[-3]); pz = __uint_as_float(w[-2]); }
            const float* d = dirs + 3 * lo;
            const float sx = rn_fma(ts, d[0], px), sy = rn_fma(ts, d[1], py), sz = rn_fma(ts, d[2], pz);   // optixPrograms.cu:194-196
            if (out.start_points) {
                pts[3 * tid] = sx;
                pts[3 * tid + 1] = sy;
                pts[3 * tid + 2] = sz;
                pts[3 * MARCH_BLOCK + 3 * tid] = ex;
                pts[3 * MARCH_BLOCK + 3 * tid + 1] = ey;
                pts[3 * MARCH_BLOCK + 3 * tid + 2] = ez;
            }
            const long long slot = base + k;
            if (slot < out.capacity) {
                if (out.t_start) out.t_start[slot] = ts;
                if (out.t_end) out.t_end[slot] = te;
                if (out.hit_cells) out.hit_cells[slot] = (int)w[5];
                if (out.seg8) store_seg8(out.seg8, slot, sx, sy, sz, ex, ey, ez);
            }
        }
        if (!out.start_points) continue;                        // packed records only: nothing to transpose
        __syncthreads();
        const int words = 3 * min(MARCH_BLOCK, block_total - k0);
        const long long w0 = 3 * (base + k0);
        const long long wcap = 3 * out.capacity;
#pragma unroll
        for (int m = 0; m < 3; m++) {
            const int wi = tid + m * MARCH_BLOCK;
            if (wi < words && w0 + wi < wcap) {
                out.start_points[w0 + wi] = pts[wi];
                out.end_points[w0 + wi] = pts[3 * MARCH_BLOCK + wi];
            }
        }
        __syncthreads();
    }
    if (!STAGED) return;
    __syncthreads();                                            // tile_s, excl, owner, the slab: free again
  }
}

static size_t onepass_smem_bytes(int cap, bool staged = false)
{
    return sizeof(uint32_t) * ((staged ? 0 : (size_t)MARCH_BLOCK * (6 * cap + 1)) + 2 * MARCH_BLOCK * 3 +
                               MARCH_BLOCK * 3 + MARCH_BLOCK + 4) + (size_t)MARCH_BLOCK * cap;
}

// staged single-pass march: persistent grid and its slab size
#define STAGED_MAX_CAP 1024                      // slot -> ray owner table: MARCH_BLOCK * cap bytes of shared memory
#define STAGED_CTAS_PER_SM 8
static long long staged_grid(long long nb)
{
    const long long g = (long long)RTX_NUM_SMS * STAGED_CTAS_PER_SM;
    return nb < g ? nb : g;
}
static size_t staged_slab_bytes(long long nb, int cap)
{
    return (size_t)staged_grid(nb) * MARCH_BLOCK * 6 * (size_t)cap * sizeof(uint32_t);
}

// Dense layout of the reference (slot = ray*arr + j, optixPrograms.cu:184), one pass.
__global__ void __launch_bounds__(MARCH_BLOCK)
march_dense_kernel(GridBits bits, int G, CamArgs cam, long long n_rays, int arr_size,
                   int* __restrict__ num_hits, float* __restrict__ view_dirs, float* __restrict__ ray_origins,
                   FillOut out)
{
    long long ray = (long long)blockIdx.x * MARCH_BLOCK + threadIdx.x;
    if (ray >= n_rays) return;
    RayGen r = raygen(cam.look_at, cam.focal, cam.aspect, cam.width, cam.height, (unsigned)(ray % cam.width),
                      (unsigned)(ray / cam.width));
    if (view_dirs) reinterpret_cast<float2*>(view_dirs)[ray] = make_float2(r.theta, r.phi);
    if (ray_origins) {
        ray_origins[3 * ray] = r.ox;
        ray_origins[3 * ray + 1] = r.oy;
        ray_origins[3 * ray + 2] = r.oz;
    }
    long long base = ray * arr_size;
    int nh = march_ray<true>(bits, G, r.ox, r.oy, r.oz, r.dx, r.dy, r.dz, arr_size,
                             [&](int j, const HitRecord& h) { store_hit(out, base + j, h); });
    num_hits[ray] = nh;
}

// main.cu:154-174 on the device (the reference's own "TODO: make this a CUDA kernel", :153)
__global__ void make_grid_kernel(int G, float* __restrict__ aabb)
{
    long long p = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    long long n = (long long)G * G * G;
    if (p >= n) return;
    int z = (int)(p % G), y = (int)((p / G) % G), x = (int)(p / ((long long)G * G));
    float L = 2.0f / (float)G;
    float* a = aabb + 6 * p;
    a[0] = rn_add(-1.0f, rn_mul((float)x, L));
    a[1] = rn_add(-1.0f, rn_mul((float)y, L));
    a[2] = rn_add(-1.0f, rn_mul((float)z, L));
    a[3] = rn_add(a[0], L);
    a[4] = rn_add(a[1], L);
    a[5] = rn_add(a[2], L);
}

// ------------------------------------------------------------------------------------ host ABI
static bool valid_res(int G) { return G >= 2 && G <= 1024 && (G & (G - 1)) == 0; }

int rtx_grid_rebuild_bricks(const rtxnerf_grid* g, cudaStream_t stream)
{
    const long long n = (long long)g->gb * g->gb * g->gb;
    brick_build_kernel<<<(unsigned)((n + 127) / 128), 128, 0, stream>>>(g->d_bits, g->G, g->gb, g->d_bricks);
    return rtx_launch_status();
}
// what the march kernels read: RTXNERF_MARCH_BRICKS=0 keeps them on the linear bits (development A/B)
static GridBits grid_bits(const rtxnerf_grid* g)
{
    static const int use_bricks = rtx_env_int("RTXNERF_MARCH_BRICKS", 1);
    return GridBits{g->d_bits, use_bricks ? g->d_bricks : nullptr, g->gb};
}

extern "C" int rtxnerf_grid_create(const uint32_t* occ_bits_host, int G, rtxnerf_grid** out)
{
    if (!occ_bits_host || !out) return RTXNERF_EINVAL;
    if (!valid_res(G)) return RTXNERF_EGRID;
    rtxnerf_grid* g = new (std::nothrow) rtxnerf_grid();
    if (!g) return RTXNERF_ENOMEM;
    g->G = G;
    g->n_words = ((size_t)G * G * G + 31) / 32;
    g->d_aabb = nullptr;
    g->d_bricks = nullptr;
    g->gb = (unsigned)((G + 3) / 4);
    cudaError_t e = cudaMalloc(&g->d_bits, g->n_words * sizeof(uint32_t));
    if (e != cudaSuccess) { delete g; return (int)e; }
    e = cudaMalloc(&g->d_bricks, (size_t)g->gb * g->gb * g->gb * sizeof(unsigned long long));
    if (e != cudaSuccess) { cudaFree(g->d_bits); delete g; return (int)e; }
    e = cudaMemcpy(g->d_bits, occ_bits_host, g->n_words * sizeof(uint32_t), cudaMemcpyHostToDevice);
    int rc = e != cudaSuccess ? (int)e : rtx_grid_rebuild_bricks(g, (cudaStream_t)0);
    if (!rc) rc = (int)cudaStreamSynchronize((cudaStream_t)0);
    if (rc) { cudaFree(g->d_bits); cudaFree(g->d_bricks); delete g; return rc; }
    *out = g;
    return RTXNERF_OK;
}

extern "C" int rtxnerf_grid_create_from_aabbs(const float* aabb6_host, int n_prims, int G, rtxnerf_grid** out)
{
    if (!aabb6_host || !out || n_prims < 0) return RTXNERF_EINVAL;
    if (!valid_res(G)) return RTXNERF_EGRID;
    std::vector<uint32_t> bits(((size_t)G * G * G + 31) / 32, 0u);
    const float L = 2.0f / (float)G;
    for (int i = 0; i < n_prims; i++) {
        const float* a = aabb6_host + 6 * (size_t)i;
        int c[3];
        for (int k = 0; k < 3; k++) {
            c[k] = (int)lrintf((a[k] + 1.0f) / L);
            if (c[k] < 0 || c[k] >= G) return RTXNERF_EINVAL;
            if (a[k] != -1.0f + (float)c[k] * L || a[k + 3] != -1.0f + (float)c[k] * L + L) return RTXNERF_EINVAL;
        }
        size_t p = ((size_t)c[0] * G + c[1]) * G + c[2];
        bits[p >> 5] |= 1u << (p & 31);
    }
    return rtxnerf_grid_create(bits.data(), G, out);
}

extern "C" int rtxnerf_grid_update(rtxnerf_grid* g, const uint32_t* occ_bits_host, void* stream)
{
    if (!g || !occ_bits_host) return RTXNERF_EINVAL;
    RTX_CUDA_TRY(cudaMemcpyAsync(g->d_bits, occ_bits_host, g->n_words * sizeof(uint32_t), cudaMemcpyHostToDevice,
                                 (cudaStream_t)stream));
    return rtx_grid_rebuild_bricks(g, (cudaStream_t)stream);
}

extern "C" int rtxnerf_grid_destroy(rtxnerf_grid* g)
{
    if (!g) return RTXNERF_OK;
    cudaFree(g->d_bits);
    cudaFree(g->d_bricks);
    if (g->d_aabb) cudaFree(g->d_aabb);
    delete g;
    return RTXNERF_OK;
}

extern "C" int rtxnerf_grid_resolution(const rtxnerf_grid* g) { return g ? g->G : RTXNERF_EINVAL; }

extern "C" int rtxnerf_grid_device_aabbs(rtxnerf_grid* g, float** d_aabb6_out)
{
    if (!g || !d_aabb6_out) return RTXNERF_EINVAL;
    if (!g->d_aabb) {
        long long n = (long long)g->G * g->G * g->G;
        RTX_CUDA_TRY(cudaMalloc(&g->d_aabb, sizeof(float) * 6 * (size_t)n));
        make_grid_kernel<<<(unsigned)((n + 255) / 256), 256>>>(g->G, g->d_aabb);
        RTX_CUDA_TRY(cudaGetLastError());
        RTX_CUDA_TRY(cudaDeviceSynchronize());
    }
    *d_aabb6_out = g->d_aabb;
    return RTXNERF_OK;
}

extern "C" size_t rtxnerf_ray_march_workspace_bytes(long long n_rays)
{
    long long nb = (n_rays + MARCH_BLOCK - 1) / MARCH_BLOCK;
    if (nb < 1) nb = 1;
    // [block_offsets nb x i64][block_sums nb x i32, padded to 8][status nb x u64][ticket u64]
    return (size_t)nb * sizeof(long long) + (((size_t)nb * sizeof(int) + 7) & ~(size_t)7) +
           (size_t)nb * sizeof(unsigned long long) + sizeof(unsigned long long) + 256;
}

static size_t march_base_workspace(long long n_rays) { return (rtxnerf_ray_march_workspace_bytes(n_rays) + 255) & ~(size_t)255; }

extern "C" size_t rtxnerf_ray_march_workspace_bytes_for(long long n_rays, int max_hits)
{
    long long nb = (n_rays + MARCH_BLOCK - 1) / MARCH_BLOCK;
    if (nb < 1) nb = 1;
    if (max_hits <= ONEPASS_MAX_CAP || max_hits > STAGED_MAX_CAP) return rtxnerf_ray_march_workspace_bytes(n_rays);
    return march_base_workspace(n_rays) + staged_slab_bytes(nb, max_hits);
}

static int march_csr_impl(const rtxnerf_grid* grid, const rtxnerf_camera* cam, long long ray_begin, long long n_rays,
                          int max_hits, const rtxnerf_hits* hits, float* seg8, void* workspace, size_t workspace_bytes,
                          void* stream);

extern "C" int rtxnerf_ray_march_csr(const rtxnerf_grid* grid, const rtxnerf_camera* cam, long long ray_begin,
                                     long long n_rays, int max_hits, const rtxnerf_hits* hits, void* workspace,
                                     size_t workspace_bytes, void* stream)
{
    return march_csr_impl(grid, cam, ray_begin, n_rays, max_hits, hits, nullptr, workspace, workspace_bytes, stream);
}

// The same march that ALSO writes every hit's entry and exit point as one 32-byte record
// d_segments8[slot] = (sx, sy, sz, ex, ey, ez, 0, 0) (32-byte aligned): what rtxnerf_render_rays_fused_seg reads with
// two 16-byte loads per sample instead of six 4-byte loads at a hit-list stride between lanes.
extern "C" int rtxnerf_ray_march_csr_seg(const rtxnerf_grid* grid, const rtxnerf_camera* cam, long long ray_begin,
                                         long long n_rays, int max_hits, const rtxnerf_hits* hits, float* d_segments8,
                                         void* workspace, size_t workspace_bytes, void* stream)
{
    if (!d_segments8 || ((uintptr_t)d_segments8 & 31u)) return RTXNERF_EINVAL;
    return march_csr_impl(grid, cam, ray_begin, n_rays, max_hits, hits, d_segments8, workspace, workspace_bytes, stream);
}

static int march_csr_impl(const rtxnerf_grid* grid, const rtxnerf_camera* cam, long long ray_begin, long long n_rays,
                          int max_hits, const rtxnerf_hits* hits, float* seg8, void* workspace, size_t workspace_bytes,
                          void* stream)
{
    if (!grid || !cam || !hits || !workspace || !cam->look_at) return RTXNERF_EINVAL;
    if (!hits->num_hits || !hits->indices || !hits->view_dirs || !hits->total_hits) return RTXNERF_EINVAL;
    // the two point arrays go together; they may be left out when the packed records are written
    if ((hits->start_points == nullptr) != (hits->end_points == nullptr)) return RTXNERF_EINVAL;
    if (!hits->start_points && !seg8) return RTXNERF_EINVAL;
    if (n_rays < 0 || ray_begin < 0 || max_hits < 1 || cam->width == 0 || cam->height == 0) return RTXNERF_EINVAL;
    if (ray_begin + n_rays > (long long)cam->width * cam->height) return RTXNERF_EINVAL;
    if (workspace_bytes < rtxnerf_ray_march_workspace_bytes(n_rays)) return RTXNERF_ECAPACITY;
    cudaStream_t s = (cudaStream_t)stream;
    if (n_rays == 0) {
        RTX_CUDA_TRY(cudaMemsetAsync(hits->total_hits, 0, sizeof(long long), s));
        return RTXNERF_OK;
    }
    long long nb = (n_rays + MARCH_BLOCK - 1) / MARCH_BLOCK;
    // workspace: [block_offsets: nb x i64][block_sums: nb x i32]
    long long* block_offsets = reinterpret_cast<long long*>(workspace);
    int* block_sums = reinterpret_cast<int*>(block_offsets + nb);
    CamArgs c{cam->look_at, cam->focal_length, cam->aspect_ratio, cam->width, cam->height};
    FillOut fo{hits->t_start, hits->t_end, hits->start_points, hits->end_points, hits->hit_cells, hits->capacity, seg8};
    // development knobs, read once per process: RTXNERF_MARCH_TWO_PASS=1 forces count -> scan -> fill,
    // RTXNERF_MARCH_FLAT=0/1 picks the loop shape (default 2: flat loop for the staged path only);
    // _ENTER_EVERY / _ENTER_MIN tune the flat loop's relaunch batching (a __constant__ per device)
    static const int two_pass = rtx_env_int("RTXNERF_MARCH_TWO_PASS", 0) == 1 ? 1 : 0;
    static const int flat = rtx_env_int("RTXNERF_MARCH_FLAT", 2);
    static const int enter_every = rtx_env_int("RTXNERF_MARCH_ENTER_EVERY", 0);
    static const int enter_min = rtx_env_int("RTXNERF_MARCH_ENTER_MIN", 0);
    static RtxPerDeviceOnce knobs_set;
    if ((enter_every || enter_min) && !rtx_device_done(knobs_set)) {
        if (enter_every) RTX_CUDA_TRY(cudaMemcpyToSymbol(c_enter_every, &enter_every, sizeof(int)));
        if (enter_min) RTX_CUDA_TRY(cudaMemcpyToSymbol(c_enter_min, &enter_min, sizeof(int)));
        rtx_device_mark(knobs_set);
    }
    if (max_hits <= ONEPASS_MAX_CAP && !two_pass) {
        unsigned long long* status = reinterpret_cast<unsigned long long*>(
            reinterpret_cast<char*>(block_sums) + (((size_t)nb * sizeof(int) + 7) & ~(size_t)7));
        unsigned long long* ticket = status + nb;
        const size_t smem = onepass_smem_bytes(max_hits);
        static RtxPerDeviceOnce smem_a, smem_b;
        RTX_CUDA_TRY((cudaError_t)rtx_set_max_smem(smem_a, march_onepass_kernel<false, false>,
                                                   (int)onepass_smem_bytes(ONEPASS_MAX_CAP)));
        RTX_CUDA_TRY((cudaError_t)rtx_set_max_smem(smem_b, march_onepass_kernel<false, true>,
                                                   (int)onepass_smem_bytes(ONEPASS_MAX_CAP)));
        RTX_CUDA_TRY(cudaMemsetAsync(status, 0, ((size_t)nb + 1) * sizeof(unsigned long long), s));
        auto kern = flat == 1 ? march_onepass_kernel<false, true> : march_onepass_kernel<false, false>;
        kern<<<(unsigned)nb, MARCH_BLOCK, smem, s>>>(
            grid_bits(grid), grid->G, c, ray_begin, n_rays, max_hits, hits->num_hits, hits->view_dirs,
            hits->ray_origins, hits->indices, fo, status, ticket, hits->total_hits, (int)nb, nullptr);
        return rtx_launch_status();
    }
    if (max_hits <= STAGED_MAX_CAP && !two_pass && max_hits > ONEPASS_MAX_CAP &&
        workspace_bytes >= rtxnerf_ray_march_workspace_bytes_for(n_rays, max_hits)) {
        // one traversal, hit records staged in the workspace (the caller sized it with ..._bytes_for)
        unsigned long long* status = reinterpret_cast<unsigned long long*>(
            reinterpret_cast<char*>(block_sums) + (((size_t)nb * sizeof(int) + 7) & ~(size_t)7));
        unsigned long long* ticket = status + nb;
        uint32_t* gstage = reinterpret_cast<uint32_t*>(reinterpret_cast<char*>(workspace) + march_base_workspace(n_rays));
        const size_t smem = onepass_smem_bytes(max_hits, true);
        static RtxPerDeviceOnce smem_c, smem_d;
        RTX_CUDA_TRY((cudaError_t)rtx_set_max_smem(smem_c, march_onepass_kernel<true, false>,
                                                   (int)onepass_smem_bytes(STAGED_MAX_CAP, true)));
        RTX_CUDA_TRY((cudaError_t)rtx_set_max_smem(smem_d, march_onepass_kernel<true, true>,
                                                   (int)onepass_smem_bytes(STAGED_MAX_CAP, true)));
        RTX_CUDA_TRY(cudaMemsetAsync(status, 0, ((size_t)nb + 1) * sizeof(unsigned long long), s));
        auto kern = flat != 0 ? march_onepass_kernel<true, true> : march_onepass_kernel<true, false>;
        kern<<<(unsigned)staged_grid(nb), MARCH_BLOCK, smem, s>>>(
            grid_bits(grid), grid->G, c, ray_begin, n_rays, max_hits, hits->num_hits, hits->view_dirs,
            hits->ray_origins, hits->indices, fo, status, ticket, hits->total_hits, (int)nb, gstage);
        return rtx_launch_status();
    }
    march_count_kernel<<<(unsigned)nb, MARCH_BLOCK, 0, s>>>(grid_bits(grid), grid->G, c, ray_begin, n_rays, max_hits,
                                                          hits->num_hits, hits->view_dirs, hits->ray_origins,
                                                          block_sums);
    block_scan_kernel<<<1, 1024, 0, s>>>(block_sums, (int)nb, block_offsets, hits->total_hits);
    march_fill_kernel<<<(unsigned)nb, MARCH_BLOCK, 0, s>>>(grid_bits(grid), grid->G, c, ray_begin, n_rays, max_hits,
                                                         hits->num_hits, block_offsets, hits->indices, fo);
    return rtx_launch_status(3);
}

extern "C" int rtxnerf_ray_march_dense(const rtxnerf_params* p, void* stream)
{
    if (!p || !p->handle || !p->look_at || !p->num_hits || !p->start_points || !p->end_points) return RTXNERF_EINVAL;
    if (p->intersection_arr_size < 1 || p->width == 0 || p->height == 0) return RTXNERF_EINVAL;
    const rtxnerf_grid* grid = reinterpret_cast<const rtxnerf_grid*>((uintptr_t)p->handle);
    long long n_rays = (long long)p->width * p->height;
    CamArgs c{p->look_at, p->focal_length, p->aspect_ratio, p->width, p->height};
    FillOut fo{p->t_start, p->t_end, p->start_points, p->end_points, nullptr, n_rays * p->intersection_arr_size};
    march_dense_kernel<<<(unsigned)((n_rays + MARCH_BLOCK - 1) / MARCH_BLOCK), MARCH_BLOCK, 0, (cudaStream_t)stream>>>(
        grid_bits(grid), grid->G, c, n_rays, p->intersection_arr_size, p->num_hits, p->viewing_direction,
        p->ray_origins, fo);
    return rtx_launch_status();
}

// ------------------------------------------------------------------------------------ no-allocation march
// rtxnerf_ray_march_csr_static: the CSR march for callers that cannot allocate (CUDA graph capture, serving
// loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
// The scratch rtxnerf_ray_march_csr takes from its caller (block offsets / block sums of the
// two-pass path, the look-back status words and the tile ticket of the one-pass path) lives in this file-scope
// device array instead: RTXNERF_MARCH_STATIC_MAX_RAYS rays at most (2^21: a 1920x1080 view fits), ~321 KB that
// every GPU which loads the library holds until exit.  The array is shared by every call on a device and keeps
// the last call's contents: the march clears the words that must start at zero (status + ticket) itself on
// `stream`, and two calls that overlap on the GPU (different streams, no dependency) race on it.
// In host code `g_march_static_scratch` is only a stand-in for the symbol: the device address comes from
// cudaGetSymbolAddress, once per device, with that device current.
#define RTXNERF_MARCH_STATIC_MAX_RAYS (1ll << 21)
#define MARCH_STATIC_BLOCKS (RTXNERF_MARCH_STATIC_MAX_RAYS / MARCH_BLOCK)
// the layout rtxnerf_ray_march_workspace_bytes() describes, for MARCH_STATIC_BLOCKS tiles
#define MARCH_STATIC_BYTES                                                                          \
    ((size_t)MARCH_STATIC_BLOCKS * sizeof(long long) + (((size_t)MARCH_STATIC_BLOCKS * sizeof(int) + 7) & ~(size_t)7) + \
     (size_t)MARCH_STATIC_BLOCKS * sizeof(unsigned long long) + sizeof(unsigned long long) + 256)
static __device__ __align__(256) unsigned char g_march_static_scratch[MARCH_STATIC_BYTES];

extern "C" long long rtxnerf_ray_march_csr_static_max_rays(void) { return RTXNERF_MARCH_STATIC_MAX_RAYS; }

extern "C" int rtxnerf_ray_march_csr_static(const rtxnerf_grid* grid, const rtxnerf_camera* cam, long long ray_begin,
                                            long long n_rays, int max_hits, const rtxnerf_hits* hits, void* stream)
{
    if (n_rays > RTXNERF_MARCH_STATIC_MAX_RAYS) return RTXNERF_ECAPACITY;   // beyond the built-in scratch
    static std::atomic<void*> scratch_of_device[256];
    int dev = 0;
    RTX_CUDA_TRY(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 256) return RTXNERF_EINVAL;
    void* ptr = scratch_of_device[dev].load(std::memory_order_acquire);
    if (!ptr) {
        RTX_CUDA_TRY(cudaGetSymbolAddress(&ptr, g_march_static_scratch));
        if (!ptr) return RTXNERF_ENOMEM;
        scratch_of_device[dev].store(ptr, std::memory_order_release);      // racing threads store the same address
    }
    // workspace_bytes = the base size: max_hits <= 32 is the one-pass kernel, larger caps take count -> scan ->
    // fill (the staged one-pass path would need a slab of hit records that does not belong in a static array)
    return rtxnerf_ray_march_csr(grid, cam, ray_begin, n_rays, max_hits, hits, ptr,
                                 rtxnerf_ray_march_workspace_bytes(RTXNERF_MARCH_STATIC_MAX_RAYS), stream);
}
