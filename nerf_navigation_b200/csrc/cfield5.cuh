// cfield5.cuh — building blocks of the warp-specialised tcgen05 field kernels (cfield5.cu: per-sample field forward /
// backward; render5.cu: fused inference renderer).  Same reference path as renderfield.cuh (NeRFNetwork.forward(x, d),
// nerf/network.py:95-124 = hash-grid encode + sigma net + trunc_exp + SH(4) + colour net + sigmoid), different machine
// mapping:
//
//   * a tile is 128 samples = the 128 TMEM lanes; thread r of a "gather" warpgroup owns sample r: it walks the 16 levels
//     (all 32 lanes of a warp on the SAME level, so neighbouring samples of a ray share corner rows in L1), interpolates
//     the 32 features in registers and writes them as a TF32 (hi, lo) pair straight into TMEM with tcgen05.st;
//   * ONE thread issues every matrix product as tcgen05.mma kind::tf32 with the A operand read from TMEM and the weight
//     tiles (pre-split hi/lo TF32, K-major no-swizzle) from shared memory: 3 products per layer (lo*hi, hi*lo, hi*hi,
//     the 3xTF32 scheme that keeps fp32 accuracy), accumulators in TMEM;
//   * a "chain" warpgroup runs the epilogues: tcgen05.ld its row of the accumulator, ReLU / masks / SH / activations,
//     split to (hi, lo) and tcgen05.st it back as the next layer's A operand.  Activations never touch shared memory or
//     registers of another thread; the LSU pipe carries only the table gathers.
//   * two tile slots (2 x 256 TMEM columns) are in flight per SM so the gathers of one tile overlap the MLP chain of the
//     other; each slot has its own MMA-issuing thread and the chain warps alternate between the slots layer by layer
//     (slot 1's epilogue runs while slot 0's next layer is on the tensor core); roles hand tiles to each other through
//     mbarriers (tcgen05.commit for MMA completion).
#pragma once
#include "fieldcore.cuh"
#include "sh.cuh"
#include "tc5.cuh"

#define C5_ROWS 128
#define C5_GATHER_WARPS 8                       // two slots x four lane quadrants
#define C5_CHAIN_WARP0 C5_GATHER_WARPS          // warps 8..11
#define C5_MMA_WARP (C5_GATHER_WARPS + 4)       // warps 12, 13: one MMA-issuing thread per slot
#define C5_THREADS ((C5_MMA_WARP + 2) * 32)     // 448
#define C5_SLOT_COLS 256
// TMEM column map of a slot
#define C5_D 0        // 64: accumulator
#define C5_FHI 64     // 32: encoder features, hi
#define C5_FLO 96     // 32: lo   (backward: columns 64..95 hold the gather threads' copy of the feature gradients)
#define C5_DF 64
#define C5_HHI 128    // 64: A operand of the hidden layers, hi
#define C5_HLO 192    // 64: lo

// ---- mbarriers of a slot
struct C5Bars {
    uint64_t f_full;    // 128 gather threads: features of the tile are in TMEM
    uint64_t f_empty;   // tcgen05.commit: the first-layer products have read the features
    uint64_t d_full;    // tcgen05.commit: a layer's accumulator is complete
    uint64_t a_ready;   // 128 chain threads: the next A operand is in TMEM (and the accumulator has been read)
    uint64_t t_free;    // 128 threads: the slot's accumulator of the finished tile has been read
    uint64_t g_full;    // backward: tcgen05.commit of the feature-gradient product
};

// ---- weight tiles: a row-major fp32 matrix element W(n, k) -> TF32 (hi, lo) tiles [N x K], K-major chunk-major
//      (tc5.cuh: 16-byte chunk (n, k/4) at (k/4 * N + n) * 16)
template <typename F>
__device__ __forceinline__ void c5_stage_tile(float *hi, float *lo, int N, int K, F w_of) {
    for (int i = threadIdx.x; i < N * K; i += blockDim.x) {
        const int n = i / K, k = i - n * K;
        uint32_t h, l;
        tc5::split_tf32(w_of(n, k), h, l);
        const int o = ((k >> 2) * N + n) * 4 + (k & 3);
        hi[o] = __uint_as_float(h);
        lo[o] = __uint_as_float(l);
    }
}

// D[128 x N] = A[128 x K] * W[N x K]^T with A = (a_hi, a_lo) in TMEM and W = (w_hi, w_lo) tiles in shared memory;
// 3xTF32, small terms first (the order fq_mma3 uses).  One thread.
template <int N, int K>
__device__ __forceinline__ void c5_issue_layer(uint32_t d, uint32_t a_hi, uint32_t a_lo, uint32_t w_hi, uint32_t w_lo) {
    const uint32_t idesc = tc5::idesc_tf32(C5_ROWS, N, tc5::MAJOR_K, tc5::MAJOR_K);
    const int ks_n = K >> 3;
#pragma unroll
    for (int ks = 0; ks < ks_n; ++ks) tc5::mma_tf32_ts(d, a_lo + ks * 8, tc5::tf32_kmajor_desc(w_hi, N, ks), idesc, ks > 0);
#pragma unroll
    for (int ks = 0; ks < ks_n; ++ks) tc5::mma_tf32_ts(d, a_hi + ks * 8, tc5::tf32_kmajor_desc(w_lo, N, ks), idesc, true);
#pragma unroll
    for (int ks = 0; ks < ks_n; ++ks) tc5::mma_tf32_ts(d, a_hi + ks * 8, tc5::tf32_kmajor_desc(w_hi, N, ks), idesc, true);
}

// ---- chain helpers (thread = TMEM lane; `lane_addr` = slot base + (32 * quadrant) << 16)
// v[16] -> (hi, lo) -> 16 columns of the A operand starting at column `col`
__device__ __forceinline__ void c5_put16(uint32_t lane_addr, int col, const float (&v)[16]) {
    uint32_t hi[16], lo[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) tc5::split_tf32(v[i], hi[i], lo[i]);
    tc5::tmem_st16(lane_addr + C5_HHI + col, hi);
    tc5::tmem_st16(lane_addr + C5_HLO + col, lo);
}
__device__ __forceinline__ void c5_get16(uint32_t lane_addr, int col, float (&v)[16]) {
    uint32_t r[16];
    tc5::tmem_ld16_nowait(lane_addr + C5_D + col, r);
    tc5::tmem_ld_wait();
#pragma unroll
    for (int i = 0; i < 16; ++i) v[i] = __uint_as_float(r[i]);
}
// accumulator (64 columns) -> ReLU -> A operand; returns the sign bits (bit j: unit j > 0).  Rolled and out of line: the
// chain warps run every line of their code once per tile, so code size is instruction-cache misses (measured: 50 % of
// their cycles were `no_instruction` stalls with everything unrolled).
template <bool MASK>
__device__ __noinline__ uint64_t c5_relu_to_a(uint32_t lane_addr) {
    uint64_t m = 0;
#pragma unroll 1
    for (int c = 0; c < 4; ++c) {
        float v[16];
        c5_get16(lane_addr, 16 * c, v);
        uint32_t bits = 0;
#pragma unroll
        for (int i = 15; i >= 0; --i) {
            v[i] = fmaxf(v[i], 0.0f);  // +0 for -0 / NaN
            // bits = bits << 1 | (v > 0): the negated integer pattern of a positive float has its sign bit set
            if (MASK) bits = __funnelshift_l(0u - __float_as_uint(v[i]), bits, 1);
        }
        if (MASK) m |= (uint64_t)bits << (16 * c);
        c5_put16(lane_addr, 16 * c, v);
    }
    return m;
}
// accumulator (64 columns) .* mask -> A operand
static __device__ __noinline__ void c5_mask_to_a(uint32_t lane_addr, uint64_t m) {
#pragma unroll 1
    for (int c = 0; c < 4; ++c) {
        float v[16];
        c5_get16(lane_addr, 16 * c, v);
        const int bits = (int)(uint32_t)(m >> (16 * c));
#pragma unroll
        for (int i = 0; i < 16; ++i) {
            int keep;  // 0 or -1: sign-extended bit i
            asm("bfe.s32 %0, %1, %2, 1;" : "=r"(keep) : "r"(bits), "r"(i));
            v[i] = __uint_as_float(__float_as_uint(v[i]) & (uint32_t)keep);
        }
        c5_put16(lane_addr, 16 * c, v);
    }
}
// publish the A operand just written / release the accumulator just read
__device__ __forceinline__ void c5_chain_arrive(uint32_t bar) {
    tc5::tmem_st_wait();
    tc5::fence_before_sync();
    tc5::mbar_arrive(bar);
}

// ---- gather helpers (thread = sample)
// The 4 interpolated features of levels 2*lp, 2*lp+1 of one sample (gridencoder.cu:75-178); all lanes of the warp walk
// the same level, the 16 gathers of the pair are issued before the first is used.
template <bool SAFE>
__device__ __forceinline__ void c5_encode_pair(const float *__restrict__ table, const FqLevel *__restrict__ lvs, int lp,
                                               const float (&x01)[3], bool valid, int align_corners, float (&feat)[4]) {
    FqRaw raw[2];
    float frac[2][3];
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        const FqLevel lv = lvs[2 * lp + h];
        uint32_t cell[3];
        fq_cell(x01, lv.scale, align_corners, cell, frac[h]);
        fq_issue<1, SAFE>(table, lv, cell, raw[h]);
    }
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
        feat[2 * h] = valid ? f0 : 0.0f;
        feat[2 * h + 1] = valid ? f1 : 0.0f;
    }
}

// The same in two halves, for a gather loop that keeps the NEXT pair's 16 gathers in flight while it interpolates the
// current one (C5_PREFETCH): issue (cell, fractions, loads into `raw`), then consume.
template <bool SAFE>
__device__ __forceinline__ void c5_pair_issue(const float *__restrict__ table, const FqLevel *__restrict__ lvs, int lp,
                                              const float (&x01)[3], int align_corners, FqRaw (&raw)[2], float (&frac)[2][3]) {
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        const FqLevel lv = lvs[2 * lp + h];
        uint32_t cell[3];
        fq_cell(x01, lv.scale, align_corners, cell, frac[h]);
        fq_issue<1, SAFE>(table, lv, cell, raw[h]);
    }
}
__device__ __forceinline__ void c5_pair_consume(const FqRaw (&raw)[2], const float (&frac)[2][3], bool valid, float (&feat)[4]) {
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
        feat[2 * h] = valid ? f0 : 0.0f;
        feat[2 * h + 1] = valid ? f1 : 0.0f;
    }
}

// 4 features -> TMEM columns [4*lp, 4*lp + 4) of the feature operand (hi at C5_FHI, lo at C5_FLO)
__device__ __forceinline__ void c5_put_features(uint32_t lane_addr, int lp, const float (&feat)[4]) {
    uint32_t hi[4], lo[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) tc5::split_tf32(feat[i], hi[i], lo[i]);
    tc5::tmem_st4(lane_addr + C5_FHI + 4 * lp, hi);
    tc5::tmem_st4(lane_addr + C5_FLO + 4 * lp, lo);
}

// Contribution of levels 2*lp, 2*lp+1 to dL/dx01 of one sample: re-gather the corners and contract the 4 feature
// gradients `df` with the encoder Jacobian (gridencoder.cu:180-223 contracted with :317-343).
template <bool SAFE>
__device__ __forceinline__ void c5_encode_pair_backward(const float *__restrict__ table, const FqLevel *__restrict__ lvs,
                                                        int lp, const float (&x01)[3], int align_corners,
                                                        const float (&df)[4], float (&gx)[3]) {
    FqRaw raw[2];
    float frac[2][3], scale[2];
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        const FqLevel lv = lvs[2 * lp + h];
        uint32_t cell[3];
        scale[h] = lv.scale;
        fq_cell(x01, lv.scale, align_corners, cell, frac[h]);
        fq_issue<1, SAFE>(table, lv, cell, raw[h]);
    }
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        float2 v[8];
        fq_select(raw[h], v);
#pragma unroll
        for (int gd = 0; gd < 3; ++gd) {
            const int d0 = gd == 0 ? 1 : 0, d1 = gd == 2 ? 1 : 2;  // the two other dims, ascending
            float a0 = 0.0f, a1 = 0.0f;
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                float w = scale[h];
                w *= (k & 1) ? frac[h][d0] : 1.0f - frac[h][d0];
                w *= (k & 2) ? frac[h][d1] : 1.0f - frac[h][d1];
                const int corner = ((k & 1) << d0) | (((k >> 1) & 1) << d1);
                a0 = fmaf(w, v[corner | (1 << gd)].x - v[corner].x, a0);
                a1 = fmaf(w, v[corner | (1 << gd)].y - v[corner].y, a1);
            }
            gx[gd] = fmaf(df[2 * h], a0, gx[gd]);
            gx[gd] = fmaf(df[2 * h + 1], a1, gx[gd]);
        }
    }
}

// ---- forward weight tiles (hi/lo), 80 KB
struct __align__(128) C5FwdWeights {
    float s0[2][64 * 32];  // sigma L1   [64 x 32]
    float s1[2][16 * 64];  // sigma L2   [16 x 64]
    float c0[2][64 * 32];  // colour L1  [64 x (16 SH | 16 sigma-net outputs, column 0 unused)]
    float c1[2][64 * 64];  // colour L2
    float c2[2][16 * 64];  // colour L3  [3 -> 16 x 64]
};

__device__ __forceinline__ void c5_stage_forward_weights(C5FwdWeights &w, const float *__restrict__ Ws0,
                                                         const float *__restrict__ Ws1, const float *__restrict__ Wc0,
                                                         const float *__restrict__ Wc1, const float *__restrict__ Wc2) {
    c5_stage_tile(w.s0[0], w.s0[1], 64, 32, [&](int n, int k) { return Ws0[n * FQ_F + k]; });
    c5_stage_tile(w.s1[0], w.s1[1], 16, 64, [&](int n, int k) { return Ws1[n * FQ_HID + k]; });
    c5_stage_tile(w.c0[0], w.c0[1], 64, 32, [&](int n, int k) {  // network.py:60: input = [SH 16 | geo_feat 15]
        return k < 16 ? Wc0[n * 31 + k] : (k == 16 ? 0.0f : Wc0[n * 31 + k - 1]);
    });
    c5_stage_tile(w.c1[0], w.c1[1], 64, 64, [&](int n, int k) { return Wc1[n * FQ_HID + k]; });
    c5_stage_tile(w.c2[0], w.c2[1], 16, 64, [&](int n, int k) { return n < 3 ? Wc2[n * FQ_HID + k] : 0.0f; });
}

