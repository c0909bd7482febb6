// hashgrid.cuh — multiresolution hash encoding (instant-ngp) shared by the forward and
// backward network kernels.  tiny-cuda-nn (the reference's encoder, main.cu:35-69,721) is
// absent from the reference archive; the published algorithm is restated here and in
// oracle/rtxnerf_oracle.c with one fixed arithmetic order so the features are bit-defined.
//
// Cost model (ncu, profiles/r01_a_*): the encoding, not the MLP, bounds the network stage —
// 128 gathers per sample against ~14 kFLOP of tensor work.  What is paid for is L1TEX
// wavefronts (distinct 128-byte lines per load instruction) and issue slots, so:
//   * the two x-neighbours of a corner pair are fetched with ONE 8-byte load whenever they
//     share an aligned pair of entries (always for even gx on hashed levels: idx ^ 1; for
//     even idx on dense levels), which removes ~1/3 of the load instructions;
//   * index arithmetic is 32-bit, hashed levels of power-of-two size use a mask, dense
//     levels take the modulo only on the (rare) wrap.
#pragma once
#include "common.cuh"

struct Levels {
    float scale[RTXNERF_MAX_LEVELS];
    uint32_t res[RTXNERF_MAX_LEVELS];
    uint32_t size[RTXNERF_MAX_LEVELS];
    uint32_t offset[RTXNERF_MAX_LEVELS];
    uint32_t mode[RTXNERF_MAX_LEVELS];   // 0 dense, 1 hashed (size = 2^k), 2 hashed (any size)
    int n_levels;
};

#define HG_PRIME_Y 2654435761u
#define HG_PRIME_Z 805459861u

// [-1,1] -> [0,1], clamped: every index below stays inside its level whatever the caller passes
// (a NaN becomes 0).  FFMA.SAT: the clamp is free.  The oracle clamps the same way.
__device__ __forceinline__ float unit_coord(float p) { return __saturatef(rn_fma(p, 0.5f, 0.5f)); }

struct LevelPoint {
    uint32_t idx[8];   // entry index of corner c = (cx) | (cy << 1) | (cz << 2) within the level
    float w[8];        // trilinear weight of corner c (fp32: what the backward scatters with)
    float wx, wy, wz;  // fractional position inside the cell
    uint32_t gx, gy, gz;   // the cell (lower corner) on this level
};

// Corner indices and weights of one level for one point (x01,y01,z01 in [0,1]).
__device__ __forceinline__ void level_point(const Levels& lv, int l, float x01, float y01, float z01, LevelPoint& p)
{
    const float scale = lv.scale[l];
    const uint32_t res = lv.res[l], size = lv.size[l], mode = lv.mode[l];
    const float px = rn_fma(scale, x01, 0.5f), py = rn_fma(scale, y01, 0.5f), pz = rn_fma(scale, z01, 0.5f);
    const float fx = floorf(px), fy = floorf(py), fz = floorf(pz);
    const float wx = rn_sub(px, fx), wy = rn_sub(py, fy), wz = rn_sub(pz, fz);
    const uint32_t gx = (uint32_t)(int)fx, gy = (uint32_t)(int)fy, gz = (uint32_t)(int)fz;
    p.gx = gx; p.gy = gy; p.gz = gz;
    p.wx = wx; p.wy = wy; p.wz = wz;
    if (mode == 0) {
        const uint32_t s2 = res * res;
        const uint32_t base = gx + gy * res + gz * s2;
        const bool wraps = base + 1u + res + s2 >= size;   // only on the far faces of the grid
#pragma unroll
        for (int c = 0; c < 8; c++) {
            uint32_t i = base + (c & 1) + ((c & 2) ? res : 0u) + ((c & 4) ? s2 : 0u);
            if (wraps) i %= size;
            p.idx[c] = i;
        }
    } else {
        const uint32_t hx0 = gx, hx1 = gx + 1u;
        const uint32_t hy0 = gy * HG_PRIME_Y, hy1 = hy0 + HG_PRIME_Y;
        const uint32_t hz0 = gz * HG_PRIME_Z, hz1 = hz0 + HG_PRIME_Z;
        const uint32_t mask = size - 1u;
#pragma unroll
        for (int c = 0; c < 8; c++) {
            const uint32_t h = ((c & 1) ? hx1 : hx0) ^ ((c & 2) ? hy1 : hy0) ^ ((c & 4) ? hz1 : hz0);
            p.idx[c] = (mode == 1) ? (h & mask) : (h % size);
        }
    }
    // w_c = ((1*wx_c)*wy_c)*wz_c in this order (1*x is exact)
    const float ux = rn_sub(1.0f, wx), uy = rn_sub(1.0f, wy), uz = rn_sub(1.0f, wz);
    const float w00 = rn_mul(ux, uy), w10 = rn_mul(wx, uy), w01 = rn_mul(ux, wy), w11 = rn_mul(wx, wy);
    p.w[0] = rn_mul(w00, uz); p.w[1] = rn_mul(w10, uz); p.w[2] = rn_mul(w01, uz); p.w[3] = rn_mul(w11, uz);
    p.w[4] = rn_mul(w00, wz); p.w[5] = rn_mul(w10, wz); p.w[6] = rn_mul(w01, wz); p.w[7] = rn_mul(w11, wz);
}

__device__ __forceinline__ uint2 ldg_u2(const void* p)
{
    uint2 v;
    asm volatile("ld.global.nc.v2.u32 {%0,%1}, [%2];" : "=r"(v.x), "=r"(v.y) : "l"(p));
    return v;
}
__device__ __forceinline__ uint32_t ldg_u1(const void* p)
{
    uint32_t v;
    asm volatile("ld.global.nc.u32 %0, [%1];" : "=r"(v) : "l"(p));
    return v;
}
__device__ __forceinline__ float2 h2_bits_to_f2(uint32_t b)
{
    return __half22float2(*reinterpret_cast<const __half2*>(&b));
}

// Trilinear blend of the 8 corner values of one level, in tiny-cuda-nn's arithmetic (grid kernel,
// N-linear interpolation): the corner weight is formed in fp32 as ((1*wx_c)*wy_c)*wz_c, ROUNDED TO THE
// PARAMETER TYPE (fp16), and accumulated with fma in the parameter type over the corners in order
// c = cx | cy << 1 | cz << 2:   result = fma((T)weight, grid_val(c), result),  T = __half.
// One HFMA2 per corner for both features (the weight half is broadcast by the operand selector)
// instead of two converts + two fp32 FMAs: 20 fewer instructions per level in a kernel bound by
// issue slots.  Round 1 accumulated in fp32 and rounded once; the accumulation
// type of the reference's dependency is the parameter type.  PARITY UNPINNED either way
// (tiny-cuda-nn is absent from the archive); the oracle restates the same sequence.
__device__ __forceinline__ uint32_t blend8(const uint32_t (&v)[8], float wx, float wy, float wz)
{
    const float ux = rn_sub(1.0f, wx), uy = rn_sub(1.0f, wy), uz = rn_sub(1.0f, wz);
    const float w00 = rn_mul(ux, uy), w10 = rn_mul(wx, uy), w01 = rn_mul(ux, wy), w11 = rn_mul(wx, wy);
    // (x0 | x1) weights of the corner pair q = cy | cz << 1
    const __half2 wq[4] = {__floats2half2_rn(rn_mul(w00, uz), rn_mul(w10, uz)),
                           __floats2half2_rn(rn_mul(w01, uz), rn_mul(w11, uz)),
                           __floats2half2_rn(rn_mul(w00, wz), rn_mul(w10, wz)),
                           __floats2half2_rn(rn_mul(w01, wz), rn_mul(w11, wz))};
    __half2 acc = __floats2half2_rn(0.0f, 0.0f);
#pragma unroll
    for (int q = 0; q < 4; q++) {
        acc = __hfma2(__low2half2(wq[q]), *reinterpret_cast<const __half2*>(&v[2 * q]), acc);
        acc = __hfma2(__high2half2(wq[q]), *reinterpret_cast<const __half2*>(&v[2 * q + 1]), acc);
    }
    return *reinterpret_cast<const uint32_t*>(&acc);
}

// Features of one level: 8 corner gathers (x-pairs merged), fp32 trilinear blend in corner
// order 0..7, one rounding to fp16 at the end.  `table` must be 8-byte aligned and the level
// offset even (checked on the host).
__device__ __forceinline__ __half2 encode_level(const __half2* __restrict__ table, const Levels& lv, int l,
                                                float x01, float y01, float z01)
{
    LevelPoint p;
    level_point(lv, l, x01, y01, z01, p);
    const uint32_t* lvl = reinterpret_cast<const uint32_t*>(table) + lv.offset[l];
    uint32_t v[8];
#pragma unroll
    for (int q = 0; q < 4; q++) {
        const uint32_t i0 = p.idx[2 * q], i1 = p.idx[2 * q + 1];
        const uint2 e = ldg_u2(lvl + (i0 & ~1u));
        const bool odd = i0 & 1u;
        v[2 * q] = odd ? e.y : e.x;
        uint32_t other = odd ? e.x : e.y;
        if ((i0 ^ i1) != 1u) other = ldg_u1(lvl + i1);
        v[2 * q + 1] = other;
    }
    const uint32_t f = blend8(v, p.wx, p.wy, p.wz);
    return *reinterpret_cast<const __half2*>(&f);
}

__device__ __forceinline__ long long row_count(const long long* d_count, long long count_scale, long long n_max)
{
    if (!d_count) return n_max;
    long long n = *d_count * count_scale;
    return n < n_max ? n : n_max;
}

// host side: validate + derive the per-level mode
static inline int levels_from(const rtxnerf_hashgrid* enc, Levels* lv)
{
    if (!enc->table || enc->n_levels < 1 || enc->n_levels > RTXNERF_MAX_LEVELS || enc->n_features != 2)
        return RTXNERF_ECONFIG;
    if (((uintptr_t)enc->table & 7u) != 0) return RTXNERF_ECONFIG;
    lv->n_levels = enc->n_levels;
    for (int l = 0; l < RTXNERF_MAX_LEVELS; l++) {
        const bool on = l < enc->n_levels;
        lv->scale[l] = on ? enc->scale[l] : 0.f;
        lv->res[l] = on ? enc->res[l] : 1u;
        lv->size[l] = on ? enc->size[l] : 2u;
        lv->offset[l] = on ? enc->offset[l] : 0u;
        lv->mode[l] = 0;
        if (!on) continue;
        const uint32_t size = enc->size[l], res = enc->res[l];
        // sizes are even so an aligned pair never straddles the end of a level
        if (size < 2 || (size & 1u) || res == 0 || (enc->offset[l] & 1u)) return RTXNERF_ECONFIG;
        const bool dense = (uint64_t)res * res * res <= (uint64_t)size;
        lv->mode[l] = dense ? 0u : (((size & (size - 1u)) == 0 && size >= 4u) ? 1u : 2u);
    }
    return RTXNERF_OK;
}

// base + zero-extended 32-bit byte offset in ONE instruction (IMAD.WIDE.U32 off, 1, base);
// left to itself the compiler rebuilds every address from 64-bit pieces (5 instructions).
__device__ __forceinline__ const char* addr_plus_u32(const char* base, uint32_t off)
{
    uint64_t r;
    asm("mad.wide.u32 %0, %1, 1, %2;" : "=l"(r) : "r"(off), "l"(reinterpret_cast<uint64_t>(base)));
    return reinterpret_cast<const char*>(r);
}

// Fast paths of encode_level for mode 0 (dense) and mode 1 (hashed, power-of-two size >= 4):
// identical arithmetic and corner order, so identical bits; straight-line code so the gathers
// of two levels can be in flight together.  All index arithmetic is done on BYTE offsets
// (entry index x 4, the table is < 4 GiB): the hash primes are pre-multiplied by 4, one LOP3
// gives the masked, pair-aligned offset, one IMAD.WIDE the address.  A dense corner block that
// wraps past the end of its level (far faces of the grid only) takes the generic path.
template <int MODE>
__device__ __forceinline__ __half2 encode_level_fast(const Levels& lv, int l, const char* __restrict__ table_bytes,
                                                     float x01, float y01, float z01)
{
    const float scale = lv.scale[l];
    const uint32_t res = lv.res[l], size = lv.size[l];
    const char* lb = table_bytes + ((size_t)lv.offset[l] << 2);
    const float px = rn_fma(scale, x01, 0.5f), py = rn_fma(scale, y01, 0.5f), pz = rn_fma(scale, z01, 0.5f);
    const float fx = floorf(px), fy = floorf(py), fz = floorf(pz);
    const float wx = rn_sub(px, fx), wy = rn_sub(py, fy), wz = rn_sub(pz, fz);
    const uint32_t gx = (uint32_t)(int)fx, gy = (uint32_t)(int)fy, gz = (uint32_t)(int)fz;
    // per corner pair q = (cy, cz): byte offset of the aligned entry pair holding corner x0,
    // whether x0 is its upper entry, and the byte offset of x1 when it lies outside the pair
    uint32_t pair_off[4], x1_off[4];
    bool odd[4];
    if (MODE == 1) {
        const uint32_t mask4 = (size - 1u) << 2, mask_pair = mask4 & ~7u;
        const uint32_t gx4 = gx << 2, gx4n = gx4 + 4u;
        const uint32_t hy0 = gy * (HG_PRIME_Y << 2), hy1 = hy0 + (HG_PRIME_Y << 2);
        const uint32_t hz0 = gz * (HG_PRIME_Z << 2), hz1 = hz0 + (HG_PRIME_Z << 2);
        const uint32_t yz4[4] = {hy0 ^ hz0, hy1 ^ hz0, hy0 ^ hz1, hy1 ^ hz1};
#pragma unroll
        for (int q = 0; q < 4; q++) {
            pair_off[q] = (gx4 ^ yz4[q]) & mask_pair;
            odd[q] = ((gx4 ^ yz4[q]) & 4u) != 0u;
            x1_off[q] = (gx4n ^ yz4[q]) & mask4;
        }
    } else {
        const uint32_t s2 = res * res;
        const uint32_t base = gx + gy * res + gz * s2;
        if (base + 1u + res + s2 >= size)
            return encode_level(reinterpret_cast<const __half2*>(table_bytes), lv, l, x01, y01, z01);
        const uint32_t base4 = base << 2, r4 = res << 2, s4 = s2 << 2;
        const uint32_t b0[4] = {base4, base4 + r4, base4 + s4, base4 + r4 + s4};
#pragma unroll
        for (int q = 0; q < 4; q++) {
            pair_off[q] = b0[q] & ~7u;
            odd[q] = (b0[q] & 4u) != 0u;
            x1_off[q] = b0[q] + 4u;
        }
    }
    uint32_t v[8];
#pragma unroll
    for (int q = 0; q < 4; q++) {
        const uint2 e = __ldg(reinterpret_cast<const uint2*>(addr_plus_u32(lb, pair_off[q])));
        v[2 * q] = odd[q] ? e.y : e.x;
        uint32_t other = odd[q] ? e.x : e.y;
        // x-neighbour outside the aligned pair: hashed level with odd gx / dense level with odd index
        const bool split = (MODE == 1) ? ((gx & 1u) != 0u) : odd[q];
        if (split) other = __ldg(reinterpret_cast<const uint32_t*>(addr_plus_u32(lb, x1_off[q])));
        v[2 * q + 1] = other;
    }
    const uint32_t f = blend8(v, wx, wy, wz);
    return *reinterpret_cast<const __half2*>(&f);
}

// one level through whichever path its mode allows
__device__ __forceinline__ __half2 encode_level_any(const Levels& lv, const char* __restrict__ table_bytes, int l,
                                                    float x01, float y01, float z01)
{
    const uint32_t m = lv.mode[l];
    if (m == 1u) return encode_level_fast<1>(lv, l, table_bytes, x01, y01, z01);
    if (m == 0u) return encode_level_fast<0>(lv, l, table_bytes, x01, y01, z01);
    return encode_level(reinterpret_cast<const __half2*>(table_bytes), lv, l, x01, y01, z01);
}

// =========================================================================================
// Lean paths (round 2): the same features as encode_level, bit for bit, in fewer instructions and with nothing
// on the conversion pipe.  What changed against encode_level_fast:
//   * floor + integer cell without FRND / F2I (quarter-rate XU pipe): s = p + 2^23 rounded DOWN
//     (FADD.RM) is 2^23 + floor(p) exactly for 0 <= p < 2^23, so floor(p) = s - 2^23, the cell is
//     the low mantissa bits of s, and p - floor(p) is exact either way.  The raw bits
//     (0x4B000000 + cell) are used as they are: hashed levels multiply by odd primes and mask
//     below bit 22, so the bias never reaches a kept bit (host check: size <= 2^22); dense
//     levels subtract one per-level constant from the linear index;
//   * ONE parity predicate per hashed level: the primes are odd, so bit 0 of
//     (gx+cx) ^ (gy+cy)*P ^ (gz+cz)*Q is (gx ^ gy ^ gz ^ cx ^ cy ^ cz) & 1;
//   * addresses are base + entry * 4 in ONE IMAD.WIDE.U32 (with a byte offset and multiplier 1 the
//     compiler turns the mad into a two-instruction 64-bit add);
//   * per-level constants come from a host-derived table in the kernel parameters (LeanLevels),
//     dense levels first, hashed power-of-two levels after them — the layout every instant-ngp
//     style configuration has; anything else keeps the generic path.
// =========================================================================================
struct LevelK {
    float scale;
    uint32_t base;    // first entry of the level inside the table
    uint32_t a;       // dense: res         hashed: size - 1
    uint32_t b;       // dense: res * res   hashed: (size - 1) & ~1
    uint32_t c;       // dense: 0x4B000000 * (1 + res + res^2) mod 2^32 (bias of the raw cell bits)
    uint32_t size;    // dense: entries of the level (wrap test)
};
struct LeanLevels {
    LevelK k[RTXNERF_MAX_LEVELS];
    int n_levels, n_dense;   // levels [0, n_dense) dense, [n_dense, n_levels) hashed with size = 2^k
    int ok;                  // 0: this configuration needs the generic path
};

static inline void lean_levels_from(const Levels& lv, LeanLevels* out)
{
    out->n_levels = lv.n_levels;
    out->ok = 1;
    int nd = 0;
    while (nd < lv.n_levels && lv.mode[nd] == 0u) nd++;
    out->n_dense = nd;
    for (int l = 0; l < RTXNERF_MAX_LEVELS; l++) {
        LevelK& k = out->k[l];
        k.scale = 0.f; k.base = 0; k.a = 0; k.b = 0; k.c = 0; k.size = 2;
        if (l >= lv.n_levels) continue;
        if (l >= nd && lv.mode[l] != 1u) out->ok = 0;
        if (!(lv.scale[l] >= 0.0f && lv.scale[l] < 4194304.0f)) out->ok = 0;
        if (lv.size[l] > (1u << 22)) out->ok = 0;
        k.scale = lv.scale[l];
        k.base = lv.offset[l];
        k.size = lv.size[l];
        if (l < nd) {
            k.a = lv.res[l];
            k.b = lv.res[l] * lv.res[l];
            k.c = 0x4B000000u * (1u + k.a + k.b);
        } else {
            k.a = lv.size[l] - 1u;
            k.b = k.a & ~1u;
        }
    }
}

// frac = p - floor(p) and the raw cell bits 0x4B000000 + floor(p), for 0 <= p < 2^23
__device__ __forceinline__ void floor_split(float p, float& frac, uint32_t& cell_bits)
{
    const float s = __fadd_rd(p, 8388608.0f);
    cell_bits = __float_as_uint(s);
    frac = rn_sub(p, rn_sub(s, 8388608.0f));
}

// address of table entry `idx` (4 bytes each).  ptxas lowers this to LEA + LEA.HI.X here (it keeps
// IMAD.WIDE.U32 only for the immediate form and prefers the integer pipe in this FMA-heavy loop;
// a register multiplier makes it split off the 64-bit add: three instructions) — measured on the
// SASS, nothing to gain by spelling it differently.
__device__ __forceinline__ const uint32_t* entry_addr(const uint32_t* base, uint32_t idx)
{
    uint64_t r;
    asm("mad.wide.u32 %0, %1, 4, %2;" : "=l"(r) : "r"(idx), "l"(reinterpret_cast<uint64_t>(base)));
    return reinterpret_cast<const uint32_t*>(r);
}

__device__ __forceinline__ uint32_t encode_hashed_lean(const LevelK& k,
                                                       const uint32_t* __restrict__ table32, float x01, float y01,
                                                       float z01)
{
    const uint32_t* lb = table32 + k.base;
    float wx, wy, wz;
    uint32_t bx, by, bz;
    floor_split(rn_fma(k.scale, x01, 0.5f), wx, bx);
    floor_split(rn_fma(k.scale, y01, 0.5f), wy, by);
    floor_split(rn_fma(k.scale, z01, 0.5f), wz, bz);
    const uint32_t mask = k.a, mask_pair = k.b;
    const uint32_t bx1 = bx + 1u;
    const uint32_t hy0 = by * HG_PRIME_Y, hy1 = hy0 + HG_PRIME_Y;
    const uint32_t hz0 = bz * HG_PRIME_Z, hz1 = hz0 + HG_PRIME_Z;
    const uint32_t yz[4] = {hy0 ^ hz0, hy1 ^ hz0, hy0 ^ hz1, hy1 ^ hz1};
    const bool odd00 = ((bx ^ by ^ bz) & 1u) != 0u;   // corner (x0, y0, z0) is the upper entry of its pair
    const bool split = (bx & 1u) != 0u;               // odd gx: x1 lies outside the pair of x0
    // all eight loads are issued before the first select: with the x1 load written over the
    // selected half of the pair (one register) ptxas issued it AFTER the pair had arrived, two
    // dependent round trips per level (ncu source page, profiles/r02_b)
    uint2 e[4];
    uint32_t x1[4];
#pragma unroll
    for (int q = 0; q < 4; q++) e[q] = __ldg(reinterpret_cast<const uint2*>(entry_addr(lb, (bx ^ yz[q]) & mask_pair)));
#pragma unroll
    for (int q = 0; q < 4; q++) {
        x1[q] = 0u;
        if (split) x1[q] = __ldg(entry_addr(lb, (bx1 ^ yz[q]) & mask));
    }
    uint32_t v[8];
#pragma unroll
    for (int q = 0; q < 4; q++) {
        const bool odd = (q == 0 || q == 3) ? odd00 : !odd00;
        v[2 * q] = odd ? e[q].y : e[q].x;
        const uint32_t other = odd ? e[q].x : e[q].y;
        v[2 * q + 1] = split ? x1[q] : other;
    }
    return blend8(v, wx, wy, wz);
}

__device__ __forceinline__ uint32_t encode_dense_lean(const LevelK& k,
                                                      const uint32_t* __restrict__ table32, float x01, float y01,
                                                      float z01)
{
    const uint32_t* lb = table32 + k.base;
    float wx, wy, wz;
    uint32_t bx, by, bz;
    floor_split(rn_fma(k.scale, x01, 0.5f), wx, bx);
    floor_split(rn_fma(k.scale, y01, 0.5f), wy, by);
    floor_split(rn_fma(k.scale, z01, 0.5f), wz, bz);
    const uint32_t res = k.a, s2 = k.b;
    const uint32_t base = bx + by * res + bz * s2 - k.c;      // gx + gy*res + gz*res^2
    const uint32_t idx[4] = {base, base + res, base + s2, base + res + s2};
    uint32_t v[8];
    if (base + 1u + res + s2 >= k.size) {
        // far faces of the grid: a corner index may pass the end of the level; idx < 2 * size,
        // so one conditional subtraction is the modulo of the generic path
#pragma unroll
        for (int c = 0; c < 8; c++) {
            uint32_t i = idx[c >> 1] + (uint32_t)(c & 1);
            if (i >= k.size) i -= k.size;
            v[c] = __ldg(entry_addr(lb, i));
        }
    } else {
        uint2 e[4];
        uint32_t x1[4];
        bool odd[4];
#pragma unroll
        for (int q = 0; q < 4; q++) {
            const uint32_t* pa = entry_addr(lb, idx[q] & ~1u);
            odd[q] = (idx[q] & 1u) != 0u;
            e[q] = __ldg(reinterpret_cast<const uint2*>(pa));
            x1[q] = 0u;
            if (odd[q]) x1[q] = __ldg(pa + 2);   // x1 opens the next pair
        }
#pragma unroll
        for (int q = 0; q < 4; q++) {
            v[2 * q] = odd[q] ? e[q].y : e[q].x;
            v[2 * q + 1] = odd[q] ? x1[q] : e[q].y;
        }
    }
    return blend8(v, wx, wy, wz);
}
