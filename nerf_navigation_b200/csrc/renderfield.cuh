// renderfield.cu — fused per-sample field for the ray paths: hash-grid encode -> sigma MLP (32-64-16) -> trunc_exp,
// SH(4) of the view direction, colour MLP ([SH 16 | geo 15] -> 64 -> 64 -> 3) -> sigmoid, in ONE kernel, and its
// backward to the sample position and direction (the pose-estimator gradient).  Reference path being fused:
// NeRFNetwork.forward(x, d) (nerf/network.py:95-124) = density (:126-143) + SHEncoder (shencoder.cu:44-70) +
// color_net (:54-68, 163-191) as called per sample by NeRFRenderer.run_cuda (renderer.py:303-305, 352-354).
// Per-sample features, hidden activations and geo_feat stay in registers; HBM sees 24 B in and 16 B out per sample
// (+ 24 B of ReLU bits when a backward follows) instead of the reference's ~2.5 KB of intermediate tensors.
//
// Same warp-cooperative layout as field.cu (fieldcore.cuh): a warp owns 16 samples, the interpolated features are the
// mma.sync m16n8k8 A fragments, every layer's accumulator fragment (columns 2t, 2t+1 of n-tile nt) is re-used as the next
// layer's A fragment (slots t, t+4 of k-tile nt) by ordering that layer's weights accordingly, 3xTF32 split products.
// One 384-thread CTA per SM shares the 76 KB of pre-split weight fragments.
#pragma once
#include "fieldcore.cuh"
#include "sh.cuh"

#ifndef RF_WARPS
#define RF_WARPS 12
#endif
#define RF_THREADS (RF_WARPS * 32)
#define RF_TILE (RF_WARPS * 16)
#define RF_GEO 15
#define RF_CIN 31  // 16 SH + 15 geo (network.py:60)

#ifndef RF_WSPLIT
#define RF_WSPLIT 0   // 1: weight fragments stored unsplit (float2) and split hi/lo in registers (measured slower: render 173 vs 118 ms); 0: pre-split float4
#endif
#if RF_WSPLIT
typedef float2 RfFrag;
#else
typedef float4 RfFrag;
#endif

struct __align__(16) RenderFieldSmem {
    RfFrag s0[1024];  // fwd: sigma L1 [nt 8][kt 4][lane]        bwd: dFeat  = dHs  W_s0   [nt 4][kt 8][lane]
    RfFrag s1[512];   // fwd: sigma L2 [nt 2][kt 8][lane]        bwd: dHs    = dO   W_s1   [nt 8][kt 2][lane]
    RfFrag c0[1024];  // fwd: colour L1 [nt 8][kt 4][lane]       bwd: dIn    = dHc1 W_c0   [nt 4][kt 8][lane]
    RfFrag c1[2048];  // fwd: colour L2 [nt 8][kt 8][lane]       bwd: dHc1   = dHc2 W_c1   [nt 8][kt 8][lane]
    RfFrag c2[256];   // fwd: colour L3 [kt 8][lane] (N = 3 -> 8) bwd: dHc2   = dRGB W_c2   [nt 8][lane] (K = 3 -> 8)
    FqLevel lv[FQ_L];
    int all_safe;
};

__device__ __forceinline__ float rf_sel4(int t, float a, float b, float c, float d) {
    return t == 0 ? a : (t == 1 ? b : (t == 2 ? c : d));
}

// ReLU the 64-wide accumulator fragment `in` (recording the sign bits) and multiply by the next layer's weights:
// out[nt] += relu(in) * W, W fragments laid out [nt NT][kt 8][lane].
template <int NT>
__device__ __forceinline__ void rf_relu_layer(const float (&in)[8][4], const RfFrag *__restrict__ frags, int lane,
                                              float (&out)[NT][4], uint32_t (&mbits)[2]) {
#pragma unroll
    for (int nt = 0; nt < NT; ++nt)
#pragma unroll
        for (int c = 0; c < 4; ++c) out[nt][c] = 0.0f;
    mbits[0] = mbits[1] = 0u;
#pragma unroll
    for (int kt = 0; kt < 8; ++kt) {
        const float c0 = in[kt][0], c1 = in[kt][1], c2 = in[kt][2], c3 = in[kt][3];
        mbits[0] |= ((c0 > 0.0f ? 1u : 0u) | (c1 > 0.0f ? 2u : 0u)) << (2 * kt);
        mbits[1] |= ((c2 > 0.0f ? 1u : 0u) | (c3 > 0.0f ? 2u : 0u)) << (2 * kt);
        uint32_t ahi[4], alo[4];
        fq_split(fmaxf(c0, 0.0f), ahi[0], alo[0]);
        fq_split(fmaxf(c2, 0.0f), ahi[1], alo[1]);
        fq_split(fmaxf(c1, 0.0f), ahi[2], alo[2]);
        fq_split(fmaxf(c3, 0.0f), ahi[3], alo[3]);
#pragma unroll
        for (int nt = 0; nt < NT; ++nt) fq_mma3(out[nt], ahi, alo, frags[(nt * 8 + kt) * 32 + lane]);
    }
}

// ------------------------------------------------------------------------------------------------ forward core
// The field of the warp's 16 samples: lane (g, t) holds rows g (h = 0) and g + 8 (h = 1).  Inputs: x01 (normalised
// position, fq_load_point), valid, the unit view direction.  Outputs in accumulator-fragment layout: o = sigma-net output
// (column 8nt+2t(+1): o[0][2h] on t == 0 is log sigma of row h), o3 = colour logits (columns 2t, 2t+1), ReLU bit masks.
template <bool SAFE>
__device__ __forceinline__ void rf_eval16(const RenderFieldSmem &s, const float *__restrict__ table, int align_corners,
                                          int lane, const float (&x01)[2][3], const bool (&valid)[2],
                                          const float (&dir)[2][3], float (&o)[2][4], float (&o3)[1][4],
                                          uint32_t (&m_sig)[2], uint32_t (&m_c1)[2], uint32_t (&m_c2)[2]) {
    const int t = lane & 3;
    FqRaw raw[2];
    float frac[2][3];
    fq_group_issue<0, SAFE>(table, s.lv[t], x01, align_corners, raw, frac);

    // ---- sigma layer 1 over the four level groups (gathers of group kt+1 in flight during the MMAs of group kt)
    float acc[8][4];
#pragma unroll
    for (int nt = 0; nt < 8; ++nt)
#pragma unroll
        for (int c = 0; c < 4; ++c) acc[nt][c] = 0.0f;
#pragma unroll 1
    for (int kt = 0; kt < 4; ++kt) {
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
        for (int nt = 0; nt < 8; ++nt) fq_mma3(acc[nt], ahi, alo, s.s0[(nt * 4 + kt) * 32 + lane]);
    }
    // ---- sigma layer 2: o = columns 8nt+2t, 8nt+2t+1 (column 0 = log sigma, 1..15 = geo_feat)
    rf_relu_layer<2>(acc, s.s1, lane, o, m_sig);

    // ---- colour layer 1: K = [SH 0-7 | SH 8-15 | o 0-7 | o 8-15]; the weight of o column 0 is zero
    float sh[2][4];  // the lane's SH slots: indices t, t+4, 8+t, 12+t of rows g, g+8
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        float v[16];
        ngp_sh4(dir[h][0], dir[h][1], dir[h][2], v);
        sh[h][0] = rf_sel4(t, v[0], v[1], v[2], v[3]);
        sh[h][1] = rf_sel4(t, v[4], v[5], v[6], v[7]);
        sh[h][2] = rf_sel4(t, v[8], v[9], v[10], v[11]);
        sh[h][3] = rf_sel4(t, v[12], v[13], v[14], v[15]);
    }
#pragma unroll
    for (int nt = 0; nt < 8; ++nt)
#pragma unroll
        for (int c = 0; c < 4; ++c) acc[nt][c] = 0.0f;
#pragma unroll
    for (int kt = 0; kt < 4; ++kt) {
        float a[4];
        if (kt < 2) {
            a[0] = sh[0][2 * kt]; a[1] = sh[1][2 * kt]; a[2] = sh[0][2 * kt + 1]; a[3] = sh[1][2 * kt + 1];
        } else {
            a[0] = o[kt - 2][0]; a[1] = o[kt - 2][2]; a[2] = o[kt - 2][1]; a[3] = o[kt - 2][3];
        }
        uint32_t ahi[4], alo[4];
#pragma unroll
        for (int c = 0; c < 4; ++c) fq_split(a[c], ahi[c], alo[c]);
#pragma unroll
        for (int nt = 0; nt < 8; ++nt) fq_mma3(acc[nt], ahi, alo, s.c0[(nt * 4 + kt) * 32 + lane]);
    }
    // ---- colour layers 2 and 3
    float acc2[8][4];
    rf_relu_layer<8>(acc, s.c1, lane, acc2, m_c1);
    rf_relu_layer<1>(acc2, s.c2, lane, o3, m_c2);
}

// Stage the forward weight fragments (pre-split hi/lo TF32 pairs) of the five layers into shared memory.
__device__ __forceinline__ void rf_stage_forward_weights(RenderFieldSmem &s, const float *__restrict__ Ws0,
                                                         const float *__restrict__ Ws1, const float *__restrict__ Wc0,
                                                         const float *__restrict__ Wc1, const float *__restrict__ Wc2) {
    const int tid = threadIdx.x;
    const int nthreads = blockDim.x;
    for (int i = tid; i < 1024; i += nthreads) {  // sigma L1: slot t -> feature 8kt+2t, t+4 -> 8kt+2t+1
        const int ln = i & 31, kt = (i >> 5) & 3, nt = i >> 7, gg = ln >> 2, tt = ln & 3;
        const float *w = Ws0 + (8 * nt + gg) * FQ_F + 8 * kt + 2 * tt;
        fq_store_frag(s.s0[i], w[0], w[1]);
    }
    for (int i = tid; i < 512; i += nthreads) {  // sigma L2
        const int ln = i & 31, kt = (i >> 5) & 7, nt = i >> 8, gg = ln >> 2, tt = ln & 3;
        const float *w = Ws1 + (8 * nt + gg) * FQ_HID + 8 * kt + 2 * tt;
        fq_store_frag(s.s1[i], w[0], w[1]);
    }
    for (int i = tid; i < 1024; i += nthreads) {  // colour L1, Wc0 [64][31]
        const int ln = i & 31, kt = (i >> 5) & 3, nt = i >> 7, gg = ln >> 2, tt = ln & 3;
        const float *w = Wc0 + (8 * nt + gg) * RF_CIN;
        float b0, b1;
        if (kt < 2) {
            b0 = w[8 * kt + tt];
            b1 = w[8 * kt + tt + 4];
        } else {
            const int col = 8 * (kt - 2) + 2 * tt;  // o column; geo index = col - 1
            b0 = col == 0 ? 0.0f : w[16 + col - 1];
            b1 = w[16 + col];
        }
        fq_store_frag(s.c0[i], b0, b1);
    }
    for (int i = tid; i < 2048; i += nthreads) {  // colour L2
        const int ln = i & 31, kt = (i >> 5) & 7, nt = i >> 8, gg = ln >> 2, tt = ln & 3;
        const float *w = Wc1 + (8 * nt + gg) * FQ_HID + 8 * kt + 2 * tt;
        fq_store_frag(s.c1[i], w[0], w[1]);
    }
    for (int i = tid; i < 256; i += nthreads) {  // colour L3, Wc2 [3][64] padded to 8 rows
        const int ln = i & 31, kt = i >> 5, gg = ln >> 2, tt = ln & 3;
        const float b0 = gg < 3 ? Wc2[gg * FQ_HID + 8 * kt + 2 * tt] : 0.0f;
        const float b1 = gg < 3 ? Wc2[gg * FQ_HID + 8 * kt + 2 * tt + 1] : 0.0f;
        fq_store_frag(s.c2[i], b0, b1);
    }
}

