// field.cu — fused field query: hash-grid encode + sigma MLP + trunc_exp, forward and backward-to-x.
// Reference path being fused: NeRFNetwork.density (nerf/network.py:126-143) = GridEncoder.forward
// (gridencoder/grid.py:140-156, gridencoder.cu:75-224) -> Linear(32,64)+ReLU -> Linear(64,16) -> trunc_exp
// (activation.py:5-17); backward = trunc_exp.backward -> Linear^T -> kernel_input_backward (gridencoder.cu:317-343).
//
// Layout: one thread per point.  The 32 encoded features (forward) / feature gradients (backward) of the CTA's
// points live in shared memory as [feature][thread] (bank-conflict free), MLP weights are staged once per CTA in
// shared memory and read as broadcast float4, so per-point activations never touch HBM.  The backward re-gathers
// the table corners (L2-resident table) instead of reading a stored dy_dx.
#include "common.cuh"

#define FQ_L 16
#define FQ_C 2
#define FQ_F 32      // L*C
#define FQ_HID 64
#define FQ_OUT 16
#define FQ_THREADS 128

struct FieldSmem {
    LevelMeta meta[FQ_L];
    float feat[FQ_F][FQ_THREADS];           // forward: features; backward: dL/dfeat
    __align__(16) float wA[FQ_F * FQ_HID];  // forward: W0^T [k][j];  backward: W0 [j][k]
    __align__(16) float wB[FQ_HID * FQ_OUT]; // forward: W1^T [j][i]; backward: W1^T [j][i]
};

__device__ __forceinline__ void fq_stage_common(FieldSmem &s, const int32_t *offsets, float S, uint32_t H,
                                                uint32_t gridtype, int align_corners) {
    if (threadIdx.x < FQ_L) s.meta[threadIdx.x] = ngp_level_meta<3>(threadIdx.x, offsets, S, H, gridtype, align_corners);
}

// 8 corner float2 values of one level for point x01; returns frac/scale for the Jacobian
__device__ __forceinline__ void fq_gather(const float (&x01)[3], const LevelMeta &m, int align_corners,
                                          const float *__restrict__ table, float2 (&v)[8], float (&frac)[3]) {
    uint32_t cell[3];
#pragma unroll
    for (int d = 0; d < 3; ++d) {
        const float p = fmaf(x01[d], m.scale, align_corners ? 0.0f : 0.5f);
        cell[d] = (uint32_t)floorf(p);
        frac[d] = p - (float)cell[d];
    }
    const float *lvl = table + (size_t)m.row_offset * FQ_C;
    uint32_t rows[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) rows[k] = ngp_row3(cell[0] + (k & 1), cell[1] + ((k >> 1) & 1), cell[2] + (k >> 2), m);
#pragma unroll
    for (int k = 0; k < 8; ++k) v[k] = ngp_ld2(lvl + (size_t)rows[k] * FQ_C);
}

__global__ void __launch_bounds__(FQ_THREADS)
k_field_forward(const float *__restrict__ x, float bound, float inv2b, const float *__restrict__ table,
                const int32_t *__restrict__ offsets, float S, uint32_t H, uint32_t gridtype, int align_corners,
                const float *__restrict__ W0, const float *__restrict__ W1, float *__restrict__ sigma,
                float *__restrict__ geo, uint64_t *__restrict__ relu_mask, int64_t B) {
    __shared__ FieldSmem s;
    const int tid = threadIdx.x;
    fq_stage_common(s, offsets, S, H, gridtype, align_corners);
    for (int i = tid; i < FQ_F * FQ_HID; i += FQ_THREADS) {  // W0 [j][k] -> wA [k][j]
        const int j = i / FQ_F, k = i - j * FQ_F;
        s.wA[k * FQ_HID + j] = W0[i];
    }
    for (int i = tid; i < FQ_HID * FQ_OUT; i += FQ_THREADS) {  // W1 [i][j] -> wB [j][i]
        const int o = i / FQ_HID, j = i - o * FQ_HID;
        s.wB[j * FQ_OUT + o] = W1[i];
    }
    __syncthreads();
    const int64_t b = (int64_t)blockIdx.x * FQ_THREADS + tid;
    if (b >= B) return;

    float x01[3];
    bool oob = false;
#pragma unroll
    for (int d = 0; d < 3; ++d) {
        x01[d] = (x[b * 3 + d] + bound) * inv2b;  // grid.py:144
        oob |= (x01[d] < 0.0f || x01[d] > 1.0f);
    }
#pragma unroll 2
    for (int l = 0; l < FQ_L; ++l) {
        float f0 = 0.0f, f1 = 0.0f;
        if (!oob) {
            float2 v[8];
            float frac[3];
            fq_gather(x01, s.meta[l], align_corners, table, v, frac);
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                float w = (k & 1) ? frac[0] : 1.0f - frac[0];
                w *= (k & 2) ? frac[1] : 1.0f - frac[1];
                w *= (k & 4) ? frac[2] : 1.0f - frac[2];
                f0 = fmaf(w, v[k].x, f0);
                f1 = fmaf(w, v[k].y, f1);
            }
        }
        s.feat[2 * l][tid] = f0;
        s.feat[2 * l + 1][tid] = f1;
    }
    // MLP: 4 chunks of 16 hidden units; each thread only touches its own feat column -> no barrier needed
    float o[FQ_OUT];
#pragma unroll
    for (int i = 0; i < FQ_OUT; ++i) o[i] = 0.0f;
    uint64_t mask = 0;
#pragma unroll 1
    for (int c = 0; c < FQ_HID / 16; ++c) {
        float h[16];
#pragma unroll
        for (int j = 0; j < 16; ++j) h[j] = 0.0f;
#pragma unroll
        for (int k = 0; k < FQ_F; ++k) {
            const float f = s.feat[k][tid];
            const float4 *w = reinterpret_cast<const float4 *>(&s.wA[k * FQ_HID + c * 16]);
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const float4 ww = w[q];
                h[4 * q + 0] = fmaf(f, ww.x, h[4 * q + 0]);
                h[4 * q + 1] = fmaf(f, ww.y, h[4 * q + 1]);
                h[4 * q + 2] = fmaf(f, ww.z, h[4 * q + 2]);
                h[4 * q + 3] = fmaf(f, ww.w, h[4 * q + 3]);
            }
        }
#pragma unroll
        for (int j = 0; j < 16; ++j) {
            const bool on = h[j] > 0.0f;
            mask |= (uint64_t)on << (c * 16 + j);
            const float hj = on ? h[j] : 0.0f;
            const float4 *w = reinterpret_cast<const float4 *>(&s.wB[(c * 16 + j) * FQ_OUT]);
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const float4 ww = w[q];
                o[4 * q + 0] = fmaf(hj, ww.x, o[4 * q + 0]);
                o[4 * q + 1] = fmaf(hj, ww.y, o[4 * q + 1]);
                o[4 * q + 2] = fmaf(hj, ww.z, o[4 * q + 2]);
                o[4 * q + 3] = fmaf(hj, ww.w, o[4 * q + 3]);
            }
        }
    }
    sigma[b] = expf(o[0]);  // trunc_exp forward, activation.py:9-11
    if (relu_mask) relu_mask[b] = mask;
    if (geo) {
        float *g = geo + b * (FQ_OUT - 1);
#pragma unroll
        for (int i = 1; i < FQ_OUT; ++i) g[i - 1] = o[i];
    }
}

template <bool HAS_GEO>
__global__ void __launch_bounds__(FQ_THREADS)
k_field_backward(const float *__restrict__ x, float bound, float inv2b, const float *__restrict__ table,
                 const int32_t *__restrict__ offsets, float S, uint32_t H, uint32_t gridtype, int align_corners,
                 const float *__restrict__ W0, const float *__restrict__ W1, const float *__restrict__ sigma,
                 const uint64_t *__restrict__ relu_mask, const float *__restrict__ grad_sigma,
                 const float *__restrict__ grad_geo, float *__restrict__ grad_x, int64_t B) {
    __shared__ FieldSmem s;
    const int tid = threadIdx.x;
    fq_stage_common(s, offsets, S, H, gridtype, align_corners);
    for (int i = tid; i < FQ_F * FQ_HID; i += FQ_THREADS) s.wA[i] = W0[i];  // [j][k]
    for (int i = tid; i < FQ_HID * FQ_OUT; i += FQ_THREADS) {              // W1 [i][j] -> wB [j][i]
        const int o = i / FQ_HID, j = i - o * FQ_HID;
        s.wB[j * FQ_OUT + o] = W1[i];
    }
    __syncthreads();
    const int64_t b = (int64_t)blockIdx.x * FQ_THREADS + tid;
    if (b >= B) return;

    // dL/do: trunc_exp backward g*exp(clamp(o0,-15,15)) == g*clamp(sigma, e^-15, e^15)  (exp is monotone)
    float go[FQ_OUT];
    {
        const float sg = sigma[b];
        const float e_lo = expf(-15.0f), e_hi = expf(15.0f);
        go[0] = grad_sigma[b] * fminf(fmaxf(sg, e_lo), e_hi);
#pragma unroll
        for (int i = 1; i < FQ_OUT; ++i) go[i] = HAS_GEO ? grad_geo[b * (FQ_OUT - 1) + i - 1] : 0.0f;
    }
    const uint64_t mask = relu_mask[b];
    // dL/dfeat[k] = sum_j mask_j * (sum_i go_i W1[i][j]) * W0[j][k], two passes of 16 features
#pragma unroll 1
    for (int kc = 0; kc < FQ_F / 16; ++kc) {
        float df[16];
#pragma unroll
        for (int k = 0; k < 16; ++k) df[k] = 0.0f;
#pragma unroll 8
        for (int j = 0; j < FQ_HID; ++j) {
            float dh;
            if (HAS_GEO) {
                dh = 0.0f;
                const float4 *w1 = reinterpret_cast<const float4 *>(&s.wB[j * FQ_OUT]);
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    const float4 ww = w1[q];
                    dh = fmaf(go[4 * q + 0], ww.x, dh);
                    dh = fmaf(go[4 * q + 1], ww.y, dh);
                    dh = fmaf(go[4 * q + 2], ww.z, dh);
                    dh = fmaf(go[4 * q + 3], ww.w, dh);
                }
            } else {
                dh = go[0] * s.wB[j * FQ_OUT];
            }
            dh = ((mask >> j) & 1ull) ? dh : 0.0f;
            const float4 *w0 = reinterpret_cast<const float4 *>(&s.wA[j * FQ_F + kc * 16]);
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const float4 ww = w0[q];
                df[4 * q + 0] = fmaf(dh, ww.x, df[4 * q + 0]);
                df[4 * q + 1] = fmaf(dh, ww.y, df[4 * q + 1]);
                df[4 * q + 2] = fmaf(dh, ww.z, df[4 * q + 2]);
                df[4 * q + 3] = fmaf(dh, ww.w, df[4 * q + 3]);
            }
        }
#pragma unroll
        for (int k = 0; k < 16; ++k) s.feat[kc * 16 + k][tid] = df[k];
    }
    // encoder input-backward with re-gathered corners (gridencoder.cu:180-223 contracted with :317-343)
    float x01[3], gx[3] = {0.0f, 0.0f, 0.0f};
    bool oob = false;
#pragma unroll
    for (int d = 0; d < 3; ++d) {
        x01[d] = (x[b * 3 + d] + bound) * inv2b;
        oob |= (x01[d] < 0.0f || x01[d] > 1.0f);
    }
    if (!oob) {
#pragma unroll 2
        for (int l = 0; l < FQ_L; ++l) {
            const LevelMeta m = s.meta[l];
            float2 v[8];
            float frac[3];
            fq_gather(x01, m, align_corners, table, v, frac);
            const float g0 = s.feat[2 * l][tid], g1 = s.feat[2 * l + 1][tid];
#pragma unroll
            for (int gd = 0; gd < 3; ++gd) {
                const int d0 = gd == 0 ? 1 : 0, d1 = gd == 2 ? 1 : 2;  // the two other dims, ascending
                float a0 = 0.0f, a1 = 0.0f;
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    float w = m.scale;
                    w *= (k & 1) ? frac[d0] : 1.0f - frac[d0];
                    w *= (k & 2) ? frac[d1] : 1.0f - frac[d1];
                    const int corner = ((k & 1) << d0) | (((k >> 1) & 1) << d1);
                    a0 = fmaf(w, v[corner | (1 << gd)].x - v[corner].x, a0);
                    a1 = fmaf(w, v[corner | (1 << gd)].y - v[corner].y, a1);
                }
                gx[gd] = fmaf(g0, a0, gx[gd]);
                gx[gd] = fmaf(g1, a1, gx[gd]);
            }
        }
    }
#pragma unroll
    for (int d = 0; d < 3; ++d) grad_x[b * 3 + d] = gx[d] * inv2b;
}

static int fq_check(const int32_t *offsets, int64_t B, const char *who) {
    (void)offsets;
    NGP_REQUIRE(B >= 0, NGP_ERR_ARG, "%s: negative B", who);
    return NGP_OK;
}

extern "C" int ngp_field_query_forward(const float *x, float bound, const float *embeddings, const int32_t *offsets,
                                       float S, uint32_t H, uint32_t gridtype, int align_corners, const float *W0,
                                       const float *W1, float *sigma, float *geo, uint64_t *relu_mask, int64_t B,
                                       ngp_stream_t stream) {
    NGP_REQUIRE(x && embeddings && offsets && W0 && W1 && sigma, NGP_ERR_ARG, "ngp_field_query_forward: null pointer");
    NGP_REQUIRE(bound > 0.0f, NGP_ERR_ARG, "ngp_field_query_forward: bound must be > 0");
    int rc = fq_check(offsets, B, "ngp_field_query_forward");
    if (rc) return rc;
    if (B == 0) return NGP_OK;
    const unsigned grid = (unsigned)ngp_div_up(B, FQ_THREADS);
    k_field_forward<<<grid, FQ_THREADS, 0, (cudaStream_t)stream>>>(
        x, bound, 1.0f / (2.0f * bound), embeddings, offsets, S, H, gridtype, align_corners, W0, W1, sigma, geo,
        relu_mask, B);
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
    if (B == 0) return NGP_OK;
    const unsigned grid = (unsigned)ngp_div_up(B, FQ_THREADS);
    const float inv2b = 1.0f / (2.0f * bound);
    if (grad_geo)
        k_field_backward<true><<<grid, FQ_THREADS, 0, (cudaStream_t)stream>>>(
            x, bound, inv2b, embeddings, offsets, S, H, gridtype, align_corners, W0, W1, sigma, relu_mask, grad_sigma,
            grad_geo, grad_x, B);
    else
        k_field_backward<false><<<grid, FQ_THREADS, 0, (cudaStream_t)stream>>>(
            x, bound, inv2b, embeddings, offsets, S, H, gridtype, align_corners, W0, W1, sigma, relu_mask, grad_sigma,
            grad_geo, grad_x, B);
    NGP_CHECK_LAUNCH("ngp_field_query_backward");
    return NGP_OK;
}
