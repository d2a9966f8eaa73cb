// fieldcore.cuh — building blocks shared by the fused field kernels (field.cu: density; renderfield.cu: density +
// colour): per-level constants, branch-free aligned-row-pair corner gathers, and the mma.sync 3xTF32 helpers.
// Reference semantics: gridencoder/src/gridencoder.cu:35-72 (index), :99-138 (cell / OOB), grid.py:144 (normalisation).
#pragma once
#include "common.cuh"

#define FQ_L 16
#define FQ_C 2
#define FQ_F 32      // L*C
#define FQ_HID 64
#define FQ_OUT 16
#define FQ_THREADS 128
#define FQ_TILE 64   // points per CTA iteration (16 = one MMA row tile per warp)

// Per-level constants of the fused kernels.  `safe` levels need no modulo: a hashed level whose row count is a power
// of two (mask = rows-1) or a dense level whose largest corner index is < rows (mask = ~0; true for every table
// GridEncoder builds, grid.py:98-124).  Anything else (tiled grids that wrap, odd hash sizes) takes the exact
// `% rows` path of gridencoder.cu:71 in a separate instantiation.
struct FqLevel {
    float scale;          // exp2f(level*S)*H - 1
    uint32_t row_offset;  // first row of the level
    uint32_t fy, fz;      // y/z multipliers: hash primes, or dense strides
    uint32_t mask;
    uint32_t hashed;
    uint32_t rows;
    uint32_t safe;
};



// ------------------------------------------------------------------------------------------------ tensor-core helpers
__device__ __forceinline__ uint32_t fq_tf32(float x) {
    uint32_t r;
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
    return r;
}
__device__ __forceinline__ void fq_split(float x, uint32_t &hi, uint32_t &lo) {
    hi = fq_tf32(x);
    lo = fq_tf32(x - __uint_as_float(hi));
}
__device__ __forceinline__ float4 fq_split_pair(float b0, float b1) {  // (hi b0, hi b1, lo b0, lo b1)
    uint32_t h0, l0, h1, l1;
    fq_split(b0, h0, l0);
    fq_split(b1, h1, l1);
    return make_float4(__uint_as_float(h0), __uint_as_float(h1), __uint_as_float(l0), __uint_as_float(l1));
}
__device__ __forceinline__ void fq_mma(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
        : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
        : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
// d += a*b with the 3xTF32 split; small terms first
__device__ __forceinline__ void fq_mma3(float (&d)[4], const uint32_t (&ahi)[4], const uint32_t (&alo)[4], float4 w) {
    const uint32_t bh0 = __float_as_uint(w.x), bh1 = __float_as_uint(w.y);
    const uint32_t bl0 = __float_as_uint(w.z), bl1 = __float_as_uint(w.w);
    fq_mma(d, alo, bh0, bh1);
    fq_mma(d, ahi, bl0, bl1);
    fq_mma(d, ahi, bh0, bh1);
}

// same product with the weight pair stored unsplit (8 B instead of 16 B of shared memory per fragment entry: half the
// LDS wavefronts on the L1 data pipe, which is what binds the colour kernels) and split in registers
__device__ __forceinline__ void fq_mma3(float (&d)[4], const uint32_t (&ahi)[4], const uint32_t (&alo)[4], float2 w) {
    uint32_t bh0, bl0, bh1, bl1;
    fq_split(w.x, bh0, bl0);
    fq_split(w.y, bh1, bl1);
    fq_mma(d, alo, bh0, bh1);
    fq_mma(d, ahi, bl0, bl1);
    fq_mma(d, ahi, bh0, bh1);
}
__device__ __forceinline__ void fq_store_frag(float4 &dst, float b0, float b1) { dst = fq_split_pair(b0, b1); }
__device__ __forceinline__ void fq_store_frag(float2 &dst, float b0, float b1) { dst = make_float2(b0, b1); }

// ------------------------------------------------------------------------------------------------ gathers
#ifndef FQ_HINT
#define FQ_HINT 0
#endif
#if FQ_HINT == 1
#define FQ_LD_COARSE "ld.global.nc.L1::evict_last"
#define FQ_LD_STREAM "ld.global.nc.L1::evict_first"
#elif FQ_HINT == 2
#define FQ_LD_COARSE "ld.global.nc.L1::evict_last"
#define FQ_LD_STREAM "ld.global.nc"
#elif FQ_HINT == 3
#define FQ_LD_COARSE "ld.global.nc"
#define FQ_LD_STREAM "ld.global.nc.L1::evict_first"
#else
#define FQ_LD_COARSE "ld.global.nc"
#define FQ_LD_STREAM "ld.global.nc"
#endif
// STREAM = 0: levels 0-3 (small, dense, L1-resident candidates); 1: levels 4-15 (no reuse inside an SM)
template <int STREAM>
__device__ __forceinline__ float4 fq_ld16(const float4 *p) {
    float4 v;
    if (STREAM)
        asm(FQ_LD_STREAM ".v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
    else
        asm(FQ_LD_COARSE ".v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
    return v;
}
template <int STREAM>
__device__ __forceinline__ float2 fq_ld8(const float2 *p) {
    float2 v;
    if (STREAM)
        asm(FQ_LD_STREAM ".v2.f32 {%0,%1}, [%2];" : "=f"(v.x), "=f"(v.y) : "l"(p));
    else
        asm(FQ_LD_COARSE ".v2.f32 {%0,%1}, [%2];" : "=f"(v.x), "=f"(v.y) : "l"(p));
    return v;
}
template <int STREAM>
__device__ __forceinline__ float2 fq_ld8_if(const float2 *p, bool pred) {  // predicated load, no branch; 0 when !pred
    float2 v;
    if (STREAM)
        asm("{\n .reg .pred p;\n setp.ne.u32 p, %3, 0;\n mov.f32 %0, 0f00000000;\n mov.f32 %1, 0f00000000;\n"
            " @p " FQ_LD_STREAM ".v2.f32 {%0,%1}, [%2];\n}" : "=&f"(v.x), "=&f"(v.y) : "l"(p), "r"((uint32_t)pred));
    else
        asm("{\n .reg .pred p;\n setp.ne.u32 p, %3, 0;\n mov.f32 %0, 0f00000000;\n mov.f32 %1, 0f00000000;\n"
            " @p " FQ_LD_COARSE ".v2.f32 {%0,%1}, [%2];\n}" : "=&f"(v.x), "=&f"(v.y) : "l"(p), "r"((uint32_t)pred));
    return v;
}

// cell / fractional position of x01 at a level (gridencoder.cu:134-138)
__device__ __forceinline__ void fq_cell(const float (&x01)[3], float scale, int align_corners, uint32_t (&cell)[3],
                                        float (&frac)[3]) {
#pragma unroll
    for (int d = 0; d < 3; ++d) {
        const float p = fmaf(x01[d], scale, align_corners ? 0.0f : 0.5f);
        cell[d] = (uint32_t)floorf(p);
        frac[d] = p - (float)cell[d];
    }
}

// Raw corner data of one (point, level): 4 aligned row pairs holding corner (x, y+dy, z+dz) (and its x-neighbour when
// both rows share the pair), 4 single rows for the x-neighbours that do not.  Issued without branches so the loads of
// several (point, level)s overlap; fq_select turns them into v[k], k = dx | dy << 1 | dz << 2.
struct FqRaw {
    float4 p[4];
    float2 q[4];
    uint32_t flags;  // bit j: row of corner x is the odd half of pair j; bit 4+j: x-neighbour is the other half
};

template <int STREAM, bool SAFE>
__device__ __forceinline__ void fq_issue(const float *__restrict__ table, const FqLevel &lv, const uint32_t (&cell)[3],
                                         FqRaw &raw) {
    const float2 *rows = reinterpret_cast<const float2 *>(table) + lv.row_offset;
    const float4 *pairs = reinterpret_cast<const float4 *>(rows);
    const uint32_t y0 = cell[1] * lv.fy, y1 = y0 + lv.fy, z0 = cell[2] * lv.fz, z1 = z0 + lv.fz;
    uint32_t yz[4];
    yz[0] = lv.hashed ? (y0 ^ z0) : (y0 + z0);
    yz[1] = lv.hashed ? (y1 ^ z0) : (y1 + z0);
    yz[2] = lv.hashed ? (y0 ^ z1) : (y0 + z1);
    yz[3] = lv.hashed ? (y1 ^ z1) : (y1 + z1);
    raw.flags = 0;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        uint32_t r0 = lv.hashed ? (cell[0] ^ yz[j]) : (cell[0] + yz[j]);
        uint32_t r1 = lv.hashed ? ((cell[0] + 1u) ^ yz[j]) : (cell[0] + 1u + yz[j]);
        if (SAFE) {
            r0 &= lv.mask;
            r1 &= lv.mask;
        } else {  // gridencoder.cu:71 `index % hashmap_size`
            r0 %= lv.rows;
            r1 %= lv.rows;
        }
        if (SAFE) {
            const bool same_pair = (r0 ^ r1) == 1u;
            raw.p[j] = fq_ld16<STREAM>(pairs + (r0 >> 1));
            raw.q[j] = fq_ld8_if<STREAM>(rows + r1, !same_pair);
            raw.flags |= ((r0 & 1u) << j) | ((same_pair ? 1u : 0u) << (4 + j));
        } else {  // no alignment assumption: two 8-byte loads
            const float2 a = fq_ld8<STREAM>(rows + r0);
            raw.p[j] = make_float4(a.x, a.y, 0.0f, 0.0f);
            raw.q[j] = fq_ld8<STREAM>(rows + r1);
        }
    }
}

__device__ __forceinline__ void fq_select(const FqRaw &raw, float2 (&v)[8]) {
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const bool odd = (raw.flags >> j) & 1u, same_pair = (raw.flags >> (4 + j)) & 1u;
        const float4 p = raw.p[j];
        v[2 * j] = odd ? make_float2(p.z, p.w) : make_float2(p.x, p.y);
        v[2 * j + 1] = same_pair ? (odd ? make_float2(p.x, p.y) : make_float2(p.z, p.w)) : raw.q[j];
    }
}

// x01 = (x + bound) / (2 bound)  (grid.py:144).  Returns false for rows past B and for points outside [0,1]^3
// (gridencoder.cu:99-123: zero features, zero gradient); those get x01 = 0.5 so their (discarded) gathers stay in range.
__device__ __forceinline__ bool fq_load_point(const float *__restrict__ x, int64_t b, int64_t B, float bound,
                                              float (&x01)[3]) {
    bool ok = b < B;
#pragma unroll
    for (int d = 0; d < 3; ++d) {
        const float xv = ok ? __ldg(x + b * 3 + d) : 0.0f;
        x01[d] = __fdiv_rn(xv + bound, 2.0f * bound);
    }
#pragma unroll
    for (int d = 0; d < 3; ++d) ok = ok && !(x01[d] < 0.0f || x01[d] > 1.0f);
    if (!ok) x01[0] = x01[1] = x01[2] = 0.5f;
    return ok;
}

__device__ __forceinline__ void fq_stage_levels(FqLevel *lvs, int *all_safe, const int32_t *offsets, float S, uint32_t H,
                                                uint32_t gridtype, int align_corners) {
    if (threadIdx.x < 32) {
        bool safe = true;
        if (threadIdx.x < FQ_L) {
            const LevelMeta m = ngp_level_meta<3>(threadIdx.x, offsets, S, H, gridtype, align_corners);
            FqLevel lv;
            lv.scale = m.scale;
            lv.row_offset = m.row_offset;
            lv.hashed = m.hashed;
            lv.rows = m.rows;
            lv.fy = m.hashed ? 2654435761u : m.m1;
            lv.fz = m.hashed ? 805459861u : m.m2;
            if (m.hashed) {
                safe = m.pow2mask != 0u;
                lv.mask = m.pow2mask;
            } else {
                const uint64_t maxc = (uint64_t)floorf(m.scale + (align_corners ? 0.0f : 0.5f)) + 1u;
                safe = maxc * (1ull + m.m1 + m.m2) < (uint64_t)m.rows;
                lv.mask = 0xffffffffu;
            }
            safe = safe && (m.row_offset & 1u) == 0u && m.rows > 0u;
            lv.safe = safe;
            lvs[threadIdx.x] = lv;
        }
        const unsigned all = __all_sync(0xffffffffu, safe);
        if (threadIdx.x == 0) *all_safe = all;
    }
}

// Issue the gathers of the lane's two rows (g, g+8) at one level; returns the fractional positions.
template <int STREAM, bool SAFE>
__device__ __forceinline__ void fq_group_issue(const float *__restrict__ table, const FqLevel &lv,
                                               const float (&x01)[2][3], int align_corners, FqRaw (&raw)[2],
                                               float (&frac)[2][3]) {
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        uint32_t cell[3];
        fq_cell(x01[h], lv.scale, align_corners, cell, frac[h]);
        fq_issue<STREAM, SAFE>(table, lv, cell, raw[h]);
    }
}

// One level group of the backward (frags = dFeat weight fragments [nt 4][kt 8][lane]): issue the gathers of level 4nt+t, compute that group's feature gradients on the
// tensor cores while they are in flight (dFeat[:, 8nt..8nt+7] = dHm W0[:, 8nt..8nt+7]), then contract them with the
// encoder Jacobian (gridencoder.cu:180-223 contracted with :317-343).
template <int STREAM, bool SAFE, typename FRAG>
__device__ __forceinline__ void fq_backward_group(const FqLevel *__restrict__ lvs, const FRAG *__restrict__ frags,
                                                  const float *__restrict__ table, int nt, int lane,
                                                  const float (&x01)[2][3], const bool (&valid)[2], int align_corners,
                                                  const float (&dhm)[8][4], float (&gx)[2][3]) {
    const FqLevel lv = lvs[4 * nt + (lane & 3)];
    FqRaw raw[2];
    float frac[2][3];
    fq_group_issue<STREAM, SAFE>(table, lv, x01, align_corners, raw, frac);
    // df = (row g c0, row g c1, row g+8 c0, row g+8 c1) of level 4nt+t
    float df[4] = {0.0f, 0.0f, 0.0f, 0.0f};
#pragma unroll
    for (int kt = 0; kt < 8; ++kt) {
        uint32_t ahi[4], alo[4];
        fq_split(dhm[kt][0], ahi[0], alo[0]);
        fq_split(dhm[kt][2], ahi[1], alo[1]);
        fq_split(dhm[kt][1], ahi[2], alo[2]);
        fq_split(dhm[kt][3], ahi[3], alo[3]);
        fq_mma3(df, ahi, alo, frags[(nt * 8 + kt) * 32 + lane]);
    }
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        float2 v[8];
        fq_select(raw[h], v);
        const float g0 = valid[h] ? df[2 * h] : 0.0f, g1 = valid[h] ? df[2 * h + 1] : 0.0f;
#pragma unroll
        for (int gd = 0; gd < 3; ++gd) {
            const int d0 = gd == 0 ? 1 : 0, d1 = gd == 2 ? 1 : 2;  // the two other dims, ascending
            float a0 = 0.0f, a1 = 0.0f;
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                float w = lv.scale;
                w *= (k & 1) ? frac[h][d0] : 1.0f - frac[h][d0];
                w *= (k & 2) ? frac[h][d1] : 1.0f - frac[h][d1];
                const int corner = ((k & 1) << d0) | (((k >> 1) & 1) << d1);
                a0 = fmaf(w, v[corner | (1 << gd)].x - v[corner].x, a0);
                a1 = fmaf(w, v[corner | (1 << gd)].y - v[corner].y, a1);
            }
            gx[h][gd] = fmaf(g0, a0, gx[h][gd]);
            gx[h][gd] = fmaf(g1, a1, gx[h][gd]);
        }
    }
}

