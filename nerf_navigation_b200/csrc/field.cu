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
#include "fieldcore.cuh"

struct __align__(16) FieldSmem {
    float4 wA[1024];  // forward: layer-1 B fragments [nt 8][kt 4][lane]; backward: dFeat B fragments [nt 4][kt 8][lane]
    float4 wB[512];   // forward: layer-2 B fragments [nt 2][kt 8][lane]; backward: dH B fragments [nt 8][kt 2][lane]
    float w1row0[FQ_HID];  // backward without dL/dgeo: W1[0][:]
    struct FqLevel lv[FQ_L];
    int all_safe;  // every level's row index needs no modulo (see FqLevel)
};

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
    fq_stage_levels(s.lv, &s.all_safe, offsets, S, H, gridtype, align_corners);
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
        fq_backward_group<0, SAFE>(s.lv, s.wA, table, 0, lane, x01, valid, align_corners, dhm, gx);
#pragma unroll 1
        for (int nt = 1; nt < 4; ++nt) fq_backward_group<1, SAFE>(s.lv, s.wA, table, nt, lane, x01, valid, align_corners, dhm, gx);

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
    fq_stage_levels(s.lv, &s.all_safe, offsets, S, H, gridtype, align_corners);
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

// default implementation: the warp-specialised tcgen05 kernels of dfield5.cu; ngp_set_field_impl(0) / NGP_B200_FIELD_IMPL=mma
// selects the mma.sync kernels above (kept as the independent fp32 cross-check).  relu_mask is opaque and its bit layout
// belongs to the implementation that wrote it: a backward must run under the same setting as its forward.
int ngp_field_impl_tc5();
int d5_field_query_forward(const float *x, float bound, const float *embeddings, const int32_t *offsets, float S, uint32_t H,
                           uint32_t gridtype, int align_corners, const float *W0, const float *W1, float *sigma, float *geo,
                           uint64_t *relu_mask, int64_t B, cudaStream_t stream);
int d5_field_query_backward(const float *x, float bound, const float *embeddings, const int32_t *offsets, float S, uint32_t H,
                            uint32_t gridtype, int align_corners, const float *W0, const float *W1, const float *sigma,
                            const uint64_t *relu_mask, const float *grad_sigma, const float *grad_geo, float *grad_x,
                            int64_t B, cudaStream_t stream);

extern "C" int ngp_field_query_forward(const float *x, float bound, const float *embeddings, const int32_t *offsets,
                                       float S, uint32_t H, uint32_t gridtype, int align_corners, const float *W0,
                                       const float *W1, float *sigma, float *geo, uint64_t *relu_mask, int64_t B,
                                       ngp_stream_t stream) {
    NGP_REQUIRE_DEVICE(x, "ngp_field_query_forward");
    if (B == 0) return NGP_OK;
    NGP_REQUIRE(x && embeddings && offsets && W0 && W1 && sigma, NGP_ERR_ARG, "ngp_field_query_forward: null pointer");
    NGP_REQUIRE(bound > 0.0f, NGP_ERR_ARG, "ngp_field_query_forward: bound must be > 0");
    int rc = fq_check(embeddings, B, "ngp_field_query_forward");
    if (rc) return rc;
    if (ngp_field_impl_tc5())
        return d5_field_query_forward(x, bound, embeddings, offsets, S, H, gridtype, align_corners, W0, W1, sigma, geo, relu_mask,
                                      B, (cudaStream_t)stream);
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
    NGP_REQUIRE_DEVICE(x, "ngp_field_query_backward");
    if (B == 0) return NGP_OK;
    NGP_REQUIRE(x && embeddings && offsets && W0 && W1 && sigma && relu_mask && grad_sigma && grad_x, NGP_ERR_ARG,
                "ngp_field_query_backward: null pointer");
    NGP_REQUIRE(bound > 0.0f, NGP_ERR_ARG, "ngp_field_query_backward: bound must be > 0");
    int rc = fq_check(embeddings, B, "ngp_field_query_backward");
    if (rc) return rc;
    if (ngp_field_impl_tc5())
        return d5_field_query_backward(x, bound, embeddings, offsets, S, H, gridtype, align_corners, W0, W1, sigma, relu_mask,
                                       grad_sigma, grad_geo, grad_x, B, (cudaStream_t)stream);
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
