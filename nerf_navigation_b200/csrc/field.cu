// field.cu — fused field query: hash-grid encode + sigma MLP + trunc_exp, forward and backward-to-x.
// Reference path being fused: NeRFNetwork.density (nerf/network.py:126-143) = GridEncoder.forward
// (gridencoder/grid.py:140-156, gridencoder.cu:75-224) -> Linear(32,64)+ReLU -> Linear(64,16) -> trunc_exp
// (activation.py:5-17); backward = trunc_exp.backward -> Linear^T -> kernel_input_backward (gridencoder.cu:317-343).
//
// Design (round-1 v2, driven by the v1 ncu capture: L1 data pipe 93 % busy, half of it MLP weight reads):
//  * Warp-cooperative layout.  A warp owns one 16-row MMA tile at a time.  Lane (g = lane/4, t = lane%4) encodes
//    rows g and g+8 of the tile at levels {4j + t : j = 0..3}.  With the K dimension of the first layer ordered as
//    (level group j) x (slot t -> feature 2*level, slot t+4 -> feature 2*level+1) the interpolated features ARE the
//    m16n8k8 A fragments: no shared-memory transpose, no per-point activation ever leaves registers.
//  * The MLP runs on the tensor cores (mma.sync m16n8k8 TF32, fp32 accumulate) with the 3xTF32 split
//    (a = a_hi + a_lo, b = b_hi + b_lo; a_lo*b_hi + a_hi*b_lo + a_hi*b_hi) so the fp32 path keeps ~2^-22 relative
//    product error (the parity bar is 1e-5).  Weights are staged once per persistent CTA in shared memory,
//    pre-split and pre-ordered as B fragments: one LDS.128 per (n-tile, k-tile) feeds both row tiles.
//  * Gathers: the two x-neighbours of a cell are the same aligned 16-byte row pair whenever their row indices differ
//    only in bit 0 (always for hashed levels with even x, for dense levels with even row) -> one LDG.128 instead of
//    two LDG.64; the odd case adds a predicated LDG.64.  L1 eviction hints (no_allocate / evict_first) on the hashed
//    levels were measured to cut the L2 hit rate to ~50 % and are NOT used.  The measured bound is one L1-miss sector per clock per SM
//    (scripts/probe_l2.py), ~91 distinct sectors per point per pass.
//  * Persistent CTAs (grid = SMs x resident CTAs) loop over 128-point tiles; warps never synchronise after staging.
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

struct __align__(16) FieldSmem {
    float4 wA[1024];  // forward: layer-1 B fragments [nt 8][kt 4][lane]; backward: dFeat B fragments [nt 4][kt 8][lane]
    float4 wB[512];   // forward: layer-2 B fragments [nt 2][kt 8][lane]; backward: dH B fragments [nt 8][kt 2][lane]
    float w1row0[FQ_HID];  // backward without dL/dgeo: W1[0][:]
    struct FqLevel lv[FQ_L];
    int all_safe;  // every level's row index needs no modulo (see FqLevel)
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

__device__ __forceinline__ void fq_stage_meta(FieldSmem &s, const int32_t *offsets, float S, uint32_t H,
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
            s.lv[threadIdx.x] = lv;
        }
        const unsigned all = __all_sync(0xffffffffu, safe);
        if (threadIdx.x == 0) s.all_safe = all;
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

// ------------------------------------------------------------------------------------------------ forward
// relu_mask layout (opaque to the caller, consumed by k_field_backward): the u64 of point b is four u16, one per
// quad lane t; bit (2*nt + c) of u16 t is hidden unit 8*nt + 2*t + c  (the accumulator-fragment columns of lane t).
template <bool SAFE>
__device__ __forceinline__ void fq_forward_tiles(const FieldSmem &s, const float *__restrict__ x, float bound,
                                                 const float *__restrict__ table, int align_corners,
                                                 float *__restrict__ sigma, float *__restrict__ geo,
                                                 uint64_t *__restrict__ relu_mask, int64_t B) {
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, g = lane >> 2, t = lane & 3;
    const int64_t ntiles = (B + FQ_TILE - 1) / FQ_TILE;
    for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        const int64_t base = tile * FQ_TILE + warp * 16;
        if (base >= B) continue;
        float x01[2][3];  // the lane's rows: h -> point base + 8*h + g
        bool valid[2];
#pragma unroll
        for (int h = 0; h < 2; ++h) valid[h] = fq_load_point(x, base + 8 * h + g, B, bound, x01[h]);

        FqRaw raw[2];
        float frac[2][3];
        fq_group_issue<0, SAFE>(table, s.lv[t], x01, align_corners, raw, frac);

        // ---- layer 1, one level group (= one k-tile) per iteration; acc[nt] = hidden 8nt+2t, 8nt+2t+1 of rows g, g+8.
        // The gathers of group kt+1 are in flight while the tensor cores work on group kt.
        float acc[8][4];
#pragma unroll
        for (int nt = 0; nt < 8; ++nt)
#pragma unroll
            for (int c = 0; c < 4; ++c) acc[nt][c] = 0.0f;
#pragma unroll 1
        for (int kt = 0; kt < 4; ++kt) {
            // A fragment (a0,a1,a2,a3) = (row g c0, row g+8 c0, row g c1, row g+8 c1) of level 4kt+t
            float a[4];
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                float2 v[8];
                fq_select(raw[h], v);
                float f0 = 0.0f, f1 = 0.0f;
#pragma unroll
                for (int k = 0; k < 8; ++k) {
                    float w = (k & 1) ? frac[h][0] : 1.0f - frac[h][0];
                    w *= (k & 2) ? frac[h][1] : 1.0f - frac[h][1];
                    w *= (k & 4) ? frac[h][2] : 1.0f - frac[h][2];
                    f0 = fmaf(w, v[k].x, f0);
                    f1 = fmaf(w, v[k].y, f1);
                }
                a[h] = valid[h] ? f0 : 0.0f;
                a[2 + h] = valid[h] ? f1 : 0.0f;
            }
            if (kt < 3) fq_group_issue<1, SAFE>(table, s.lv[4 * (kt + 1) + t], x01, align_corners, raw, frac);
            uint32_t ahi[4], alo[4];
#pragma unroll
            for (int c = 0; c < 4; ++c) fq_split(a[c], ahi[c], alo[c]);
#pragma unroll
            for (int nt = 0; nt < 8; ++nt) fq_mma3(acc[nt], ahi, alo, s.wA[(nt * 4 + kt) * 32 + lane]);
        }

        // ---- ReLU (+ mask) and layer 2; accumulator columns (2t, 2t+1) of n-tile nt are k slots (t, t+4) of k-tile nt
        float out[2][4];
#pragma unroll
        for (int nt = 0; nt < 2; ++nt)
#pragma unroll
            for (int c = 0; c < 4; ++c) out[nt][c] = 0.0f;
        uint32_t mbits[2] = {0u, 0u};
#pragma unroll
        for (int kt = 0; kt < 8; ++kt) {
            const float c0 = acc[kt][0], c1 = acc[kt][1], c2 = acc[kt][2], c3 = acc[kt][3];
            mbits[0] |= ((c0 > 0.0f ? 1u : 0u) | (c1 > 0.0f ? 2u : 0u)) << (2 * kt);
            mbits[1] |= ((c2 > 0.0f ? 1u : 0u) | (c3 > 0.0f ? 2u : 0u)) << (2 * kt);
            uint32_t ahi[4], alo[4];
            fq_split(fmaxf(c0, 0.0f), ahi[0], alo[0]);
            fq_split(fmaxf(c2, 0.0f), ahi[1], alo[1]);
            fq_split(fmaxf(c1, 0.0f), ahi[2], alo[2]);
            fq_split(fmaxf(c3, 0.0f), ahi[3], alo[3]);
#pragma unroll
            for (int nt = 0; nt < 2; ++nt) fq_mma3(out[nt], ahi, alo, s.wB[(nt * 8 + kt) * 32 + lane]);
        }

        // ---- outputs: out[nt] = columns 8nt+2t, 8nt+2t+1 of rows g (c0,c1) and g+8 (c2,c3)
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const int64_t b = base + 8 * h + g;
            if (b >= B) continue;
            if (relu_mask) reinterpret_cast<uint16_t *>(relu_mask)[b * 4 + t] = (uint16_t)mbits[h];
            if (t == 0) sigma[b] = expf(out[0][2 * h]);  // trunc_exp forward, activation.py:9-11
            if (geo) {
                float *gp = geo + b * (FQ_OUT - 1);
#pragma unroll
                for (int nt = 0; nt < 2; ++nt) {
                    const int col = 8 * nt + 2 * t;
                    if (col > 0) gp[col - 1] = out[nt][2 * h];
                    gp[col] = out[nt][2 * h + 1];
                }
            }
        }
    }
}

#ifndef FQ_MINB_FWD
#define FQ_MINB_FWD 3
#endif
#ifndef FQ_MINB_BWD
#define FQ_MINB_BWD 3
#endif
__global__ void __launch_bounds__(FQ_THREADS, FQ_MINB_FWD)
k_field_forward(const float *__restrict__ x, float bound, const float *__restrict__ table,
                const int32_t *__restrict__ offsets, float S, uint32_t H, uint32_t gridtype, int align_corners,
                const float *__restrict__ W0, const float *__restrict__ W1, float *__restrict__ sigma,
                float *__restrict__ geo, uint64_t *__restrict__ relu_mask, int64_t B) {
    __shared__ FieldSmem s;
    const int tid = threadIdx.x;
    fq_stage_meta(s, offsets, S, H, gridtype, align_corners);
    // layer 1: B[k][n] = W0[n][f(k)], k-tile kt slot t -> feature 8kt+2t, slot t+4 -> feature 8kt+2t+1
    for (int i = tid; i < 1024; i += FQ_THREADS) {
        const int ln = i & 31, kt = (i >> 5) & 3, nt = i >> 7, gg = ln >> 2, tt = ln & 3;
        const float *w = W0 + (8 * nt + gg) * FQ_F + 8 * kt + 2 * tt;
        s.wA[i] = fq_split_pair(w[0], w[1]);
    }
    // layer 2: B[k][n] = W1[n][h(k)], k-tile kt slot t -> hidden 8kt+2t, slot t+4 -> hidden 8kt+2t+1
    for (int i = tid; i < 512; i += FQ_THREADS) {
        const int ln = i & 31, kt = (i >> 5) & 7, nt = i >> 8, gg = ln >> 2, tt = ln & 3;
        const float *w = W1 + (8 * nt + gg) * FQ_HID + 8 * kt + 2 * tt;
        s.wB[i] = fq_split_pair(w[0], w[1]);
    }
    __syncthreads();
    if (s.all_safe) fq_forward_tiles<true>(s, x, bound, table, align_corners, sigma, geo, relu_mask, B);
    else fq_forward_tiles<false>(s, x, bound, table, align_corners, sigma, geo, relu_mask, B);
}


// ------------------------------------------------------------------------------------------------ backward to x
// One level group of the backward: issue the gathers of level 4nt+t, compute that group's feature gradients on the
// tensor cores while they are in flight (dFeat[:, 8nt..8nt+7] = dHm W0[:, 8nt..8nt+7]), then contract them with the
// encoder Jacobian (gridencoder.cu:180-223 contracted with :317-343).
template <int STREAM, bool SAFE>
__device__ __forceinline__ void fq_backward_group(const FieldSmem &s, const float *__restrict__ table, int nt, int lane,
                                                  const float (&x01)[2][3], const bool (&valid)[2], int align_corners,
                                                  const float (&dhm)[8][4], float (&gx)[2][3]) {
    const FqLevel lv = s.lv[4 * nt + (lane & 3)];
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
        fq_mma3(df, ahi, alo, s.wA[(nt * 8 + kt) * 32 + lane]);
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

template <bool HAS_GEO, bool SAFE>
__device__ __forceinline__ void fq_backward_tiles(const FieldSmem &s, const float *__restrict__ x, float bound,
                                                  const float *__restrict__ table, int align_corners,
                                                  const float *__restrict__ sigma,
                                                  const uint64_t *__restrict__ relu_mask,
                                                  const float *__restrict__ grad_sigma,
                                                  const float *__restrict__ grad_geo, float *__restrict__ grad_x,
                                                  int64_t B) {
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, g = lane >> 2, t = lane & 3;
    const float e_lo = expf(-15.0f), e_hi = expf(15.0f);
    const float den = 2.0f * bound;
    const int64_t ntiles = (B + FQ_TILE - 1) / FQ_TILE;
    for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        const int64_t base = tile * FQ_TILE + warp * 16;
        if (base >= B) continue;
        float x01[2][3];
        bool valid[2];
        float go0[2];  // dL/do[0] = dL/dsigma * exp(clamp(o0,-15,15)) = g * clamp(sigma, e^-15, e^15)  (activation.py:13-17)
        uint32_t mbits[2];
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const int64_t b = base + 8 * h + g;
            valid[h] = fq_load_point(x, b, B, bound, x01[h]);
            const bool inb = b < B;
            go0[h] = inb ? __ldg(grad_sigma + b) * fminf(fmaxf(__ldg(sigma + b), e_lo), e_hi) : 0.0f;
            mbits[h] = inb ? (uint32_t)__ldg(reinterpret_cast<const uint16_t *>(relu_mask) + b * 4 + t) : 0u;
        }

        // ---- dH . relu' in accumulator-fragment layout: dhm[nt] = units 8nt+2t, 8nt+2t+1 of rows g (c0,c1), g+8 (c2,c3)
        float dhm[8][4];
        if (HAS_GEO) {
#pragma unroll
            for (int nt = 0; nt < 8; ++nt)
#pragma unroll
                for (int c = 0; c < 4; ++c) dhm[nt][c] = 0.0f;
#pragma unroll
            for (int kt = 0; kt < 2; ++kt) {
                uint32_t ahi[4], alo[4];
#pragma unroll
                for (int c = 0; c < 4; ++c) {  // a_c: row g + 8*(c&1), output column 8kt + t + 4*(c>>1)
                    const int64_t b = base + 8 * (c & 1) + g;
                    const int col = 8 * kt + t + 4 * (c >> 1);
                    float v = 0.0f;
                    if (b < B) v = (col == 0) ? go0[c & 1] : __ldg(grad_geo + b * (FQ_OUT - 1) + (col > 0 ? col - 1 : 0));
                    fq_split(v, ahi[c], alo[c]);
                }
#pragma unroll
                for (int nt = 0; nt < 8; ++nt) fq_mma3(dhm[nt], ahi, alo, s.wB[(nt * 2 + kt) * 32 + lane]);
            }
        } else {
#pragma unroll
            for (int nt = 0; nt < 8; ++nt) {
                const float2 w = *reinterpret_cast<const float2 *>(&s.w1row0[8 * nt + 2 * t]);
                dhm[nt][0] = go0[0] * w.x;
                dhm[nt][1] = go0[0] * w.y;
                dhm[nt][2] = go0[1] * w.x;
                dhm[nt][3] = go0[1] * w.y;
            }
        }
#pragma unroll
        for (int nt = 0; nt < 8; ++nt) {
            const uint32_t m0 = mbits[0] >> (2 * nt), m1 = mbits[1] >> (2 * nt);
            dhm[nt][0] = (m0 & 1u) ? dhm[nt][0] : 0.0f;
            dhm[nt][1] = (m0 & 2u) ? dhm[nt][1] : 0.0f;
            dhm[nt][2] = (m1 & 1u) ? dhm[nt][2] : 0.0f;
            dhm[nt][3] = (m1 & 2u) ? dhm[nt][3] : 0.0f;
        }

        float gx[2][3];
#pragma unroll
        for (int h = 0; h < 2; ++h) gx[h][0] = gx[h][1] = gx[h][2] = 0.0f;
        fq_backward_group<0, SAFE>(s, table, 0, lane, x01, valid, align_corners, dhm, gx);
#pragma unroll 1
        for (int nt = 1; nt < 4; ++nt) fq_backward_group<1, SAFE>(s, table, nt, lane, x01, valid, align_corners, dhm, gx);

        // sum the four level groups held by the quad; lane t < 3 stores component t
#pragma unroll
        for (int h = 0; h < 2; ++h) {
#pragma unroll
            for (int d = 0; d < 3; ++d) {
                gx[h][d] += __shfl_xor_sync(0xffffffffu, gx[h][d], 1);
                gx[h][d] += __shfl_xor_sync(0xffffffffu, gx[h][d], 2);
            }
            const int64_t b = base + 8 * h + g;
            const float mine = t == 0 ? gx[h][0] : (t == 1 ? gx[h][1] : gx[h][2]);
            if (b < B && t < 3) grad_x[b * 3 + t] = __fdiv_rn(mine, den);
        }
    }
}

template <bool HAS_GEO>
__global__ void __launch_bounds__(FQ_THREADS, FQ_MINB_BWD)
k_field_backward(const float *__restrict__ x, float bound, const float *__restrict__ table,
                 const int32_t *__restrict__ offsets, float S, uint32_t H, uint32_t gridtype, int align_corners,
                 const float *__restrict__ W0, const float *__restrict__ W1, const float *__restrict__ sigma,
                 const uint64_t *__restrict__ relu_mask, const float *__restrict__ grad_sigma,
                 const float *__restrict__ grad_geo, float *__restrict__ grad_x, int64_t B) {
    __shared__ FieldSmem s;
    const int tid = threadIdx.x;
    fq_stage_meta(s, offsets, S, H, gridtype, align_corners);
    // dFeat = dH W0: B[k][n] = W0[h(k)][n], k-tile kt slot t -> hidden 8kt+2t, slot t+4 -> hidden 8kt+2t+1; n natural
    for (int i = tid; i < 1024; i += FQ_THREADS) {
        const int ln = i & 31, kt = (i >> 5) & 7, nt = i >> 8, gg = ln >> 2, tt = ln & 3;
        const int n = 8 * nt + gg;
        s.wA[i] = fq_split_pair(W0[(8 * kt + 2 * tt) * FQ_F + n], W0[(8 * kt + 2 * tt + 1) * FQ_F + n]);
    }
    if (HAS_GEO) {
        // dH = dO W1: B[k][n] = W1[k][n], natural order, K = 16 outputs
        for (int i = tid; i < 512; i += FQ_THREADS) {
            const int ln = i & 31, kt = (i >> 5) & 1, nt = i >> 6, gg = ln >> 2, tt = ln & 3;
            const int n = 8 * nt + gg;
            s.wB[i] = fq_split_pair(W1[(8 * kt + tt) * FQ_HID + n], W1[(8 * kt + tt + 4) * FQ_HID + n]);
        }
    } else {
        if (tid < FQ_HID) s.w1row0[tid] = W1[tid];
    }
    __syncthreads();
    if (s.all_safe)
        fq_backward_tiles<HAS_GEO, true>(s, x, bound, table, align_corners, sigma, relu_mask, grad_sigma, grad_geo, grad_x, B);
    else
        fq_backward_tiles<HAS_GEO, false>(s, x, bound, table, align_corners, sigma, relu_mask, grad_sigma, grad_geo, grad_x, B);
}
// ------------------------------------------------------------------------------------------------ C ABI
template <typename K>
static int fq_grid(K kernel, int64_t B) {
    int dev = 0, sms = NGP_SM_COUNT, per_sm = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, FQ_THREADS, 0) != cudaSuccess || per_sm < 1) per_sm = 1;
    const int64_t tiles = ngp_div_up(B, FQ_TILE), resident = (int64_t)sms * per_sm;
    return (int)(tiles < resident ? tiles : resident);
}

static int fq_check(const float *embeddings, int64_t B, const char *who) {
    NGP_REQUIRE(B >= 0, NGP_ERR_ARG, "%s: negative B", who);
    NGP_REQUIRE(((uintptr_t)embeddings & 15u) == 0, NGP_ERR_ARG, "%s: embeddings must be 16-byte aligned", who);
    return NGP_OK;
}

extern "C" int ngp_field_query_forward(const float *x, float bound, const float *embeddings, const int32_t *offsets,
                                       float S, uint32_t H, uint32_t gridtype, int align_corners, const float *W0,
                                       const float *W1, float *sigma, float *geo, uint64_t *relu_mask, int64_t B,
                                       ngp_stream_t stream) {
    NGP_REQUIRE(x && embeddings && offsets && W0 && W1 && sigma, NGP_ERR_ARG, "ngp_field_query_forward: null pointer");
    NGP_REQUIRE(bound > 0.0f, NGP_ERR_ARG, "ngp_field_query_forward: bound must be > 0");
    int rc = fq_check(embeddings, B, "ngp_field_query_forward");
    if (rc) return rc;
    if (B == 0) return NGP_OK;
    k_field_forward<<<fq_grid(k_field_forward, B), FQ_THREADS, 0, (cudaStream_t)stream>>>(
        x, bound, embeddings, offsets, S, H, gridtype, align_corners, W0, W1, sigma, geo, relu_mask, B);
    NGP_CHECK_LAUNCH("ngp_field_query_forward");
    return NGP_OK;
}

extern "C" int ngp_field_query_backward(const float *x, float bound, const float *embeddings, const int32_t *offsets,
                                        float S, uint32_t H, uint32_t gridtype, int align_corners, const float *W0,
                                        const float *W1, const float *sigma, const uint64_t *relu_mask,
                                        const float *grad_sigma, const float *grad_geo, float *grad_x, int64_t B,
                                        ngp_stream_t stream) {
    NGP_REQUIRE(x && embeddings && offsets && W0 && W1 && sigma && relu_mask && grad_sigma && grad_x, NGP_ERR_ARG,
                "ngp_field_query_backward: null pointer");
    NGP_REQUIRE(bound > 0.0f, NGP_ERR_ARG, "ngp_field_query_backward: bound must be > 0");
    int rc = fq_check(embeddings, B, "ngp_field_query_backward");
    if (rc) return rc;
    if (B == 0) return NGP_OK;
    if (grad_geo)
        k_field_backward<true><<<fq_grid(k_field_backward<true>, B), FQ_THREADS, 0, (cudaStream_t)stream>>>(
            x, bound, embeddings, offsets, S, H, gridtype, align_corners, W0, W1, sigma, relu_mask, grad_sigma,
            grad_geo, grad_x, B);
    else
        k_field_backward<false><<<fq_grid(k_field_backward<false>, B), FQ_THREADS, 0, (cudaStream_t)stream>>>(
            x, bound, embeddings, offsets, S, H, gridtype, align_corners, W0, W1, sigma, relu_mask, grad_sigma,
            grad_geo, grad_x, B);
    NGP_CHECK_LAUNCH("ngp_field_query_backward");
    return NGP_OK;
}
