// fieldvg.cu — density value AND gradient in ONE pass: sigma(x) and u * d sigma / d x for every point, where the upstream
// factor u_i = a_i + b_i * sigma_i is known per point (u = 1: the plain Jacobian; a = 0, b = 2 c_i: the planner's
// collision cost sum_i c_i sigma_i^2, nav/quad_plot.py:224-250; b = 2/B: mean(sigma^2)).
// Reference path: NeRFNetwork.density (nerf/network.py:126-143) followed by autograd's backward through trunc_exp, the two
// Linear layers and the hash-grid encoder (gridencoder.cu:180-223, 317-343) — two kernels / two full gather passes in
// field.cu.  Because sigma_i depends on x_i only, the backward needs nothing the forward has not already fetched: the 8
// corner rows of a (point, level) cell are kept in SHARED MEMORY between the two halves (levels 4-15: 12 x 64 B per
// point, 144 KB per 192-point CTA), so the L2-resident table is gathered once (plus the four coarse levels, which mostly
// hit L1).  The gather traffic of a "sample fwd+bwd" drops from 2048 B to 1024 + 256 B.
// Same warp-cooperative layout as field.cu: lane (g, t) = rows g, g+8 x level 4j + t; features are the mma.sync A
// fragments; 3xTF32 split products; one persistent 384-thread CTA per SM.
#include "fieldcore.cuh"

#ifndef VG_PIPE
#define VG_PIPE 0      // 1: cache all 16 levels and prefetch the next tile's first level group during the backward (measured slower: 3.7 vs 2.7 ms)
#endif
#if VG_PIPE
#define VG_WARPS 11
#define VG_CACHED 16
#define VG_G0 0        // first cached level group
#else
#define VG_WARPS 12
#define VG_CACHED 12   // levels 4..15 (level groups 1..3); group 0 is re-gathered (mostly L1 hits)
#define VG_G0 1
#endif
#define VG_THREADS (VG_WARPS * 32)
#define VG_TILE (VG_WARPS * 16)

struct __align__(16) FieldVGSmem {
    float4 wF1[1024];      // forward layer 1 B fragments [nt 8][kt 4][lane]
    float4 wF2[512];       // forward layer 2 B fragments [nt 2][kt 8][lane]
    float4 wB1[1024];      // backward dFeat = dH W0 B fragments [nt 4][kt 8][lane]
    float w1row0[FQ_HID];  // W1[0][:]  (d o_0 / d hidden)
    FqLevel lv[FQ_L];
    int all_safe;
    // corner cache: [warp][corner pair 4][row 16][cached level 12] float4 = (v[2p].x, v[2p].y, v[2p+1].x, v[2p+1].y);
    // level fastest, so a quarter warp (2 rows x 4 levels) touches 32 distinct banks
    float4 cache[VG_WARPS][4][16][VG_CACHED];
};

// dFeat of level group nt for the lane's level (4nt + t): df = (row g c0, row g c1, row g+8 c0, row g+8 c1)
__device__ __forceinline__ void vg_dfeat(const float4 *__restrict__ frags, int nt, int lane, const float (&dhm)[8][4],
                                         float (&df)[4]) {
    df[0] = df[1] = df[2] = df[3] = 0.0f;
#pragma unroll
    for (int kt = 0; kt < 8; ++kt) {
        uint32_t ahi[4], alo[4];
        fq_split(dhm[kt][0], ahi[0], alo[0]);
        fq_split(dhm[kt][2], ahi[1], alo[1]);
        fq_split(dhm[kt][1], ahi[2], alo[2]);
        fq_split(dhm[kt][3], ahi[3], alo[3]);
        fq_mma3(df, ahi, alo, frags[(nt * 8 + kt) * 32 + lane]);
    }
}

// gx += (g0, g1) . d(feat)/dx of one (row, level): gridencoder.cu:180-223 contracted with :317-343
__device__ __forceinline__ void vg_contract(const float2 (&v)[8], const float (&frac)[3], float scale, float g0, float g1,
                                            float (&gx)[3]) {
#pragma unroll
    for (int gd = 0; gd < 3; ++gd) {
        const int d0 = gd == 0 ? 1 : 0, d1 = gd == 2 ? 1 : 2;  // the two other dims, ascending
        float a0 = 0.0f, a1 = 0.0f;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            float w = scale;
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

template <bool SAFE>
__device__ __forceinline__ void vg_tiles(FieldVGSmem &s, const float *__restrict__ x, float bound,
                                         const float *__restrict__ table, int align_corners,
                                         const float *__restrict__ coef_a, const float *__restrict__ coef_b, float a0,
                                         float b0, float *__restrict__ sigma, float *__restrict__ geo,
                                         float *__restrict__ grad_x, int64_t B) {
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, g = lane >> 2, t = lane & 3;
    const float e_lo = expf(-15.0f), e_hi = expf(15.0f);
    const float den = 2.0f * bound;
    float4 (*cache)[16][VG_CACHED] = s.cache[warp];
    const int64_t ntiles = (B + VG_TILE - 1) / VG_TILE;
    float x01[2][3];
    bool valid[2];
    FqRaw raw[2];
    float frac[2][3];
#if VG_PIPE
    {   // prologue: the first tile's points and first level group
        const int64_t base0 = (int64_t)blockIdx.x * VG_TILE + warp * 16;
#pragma unroll
        for (int h = 0; h < 2; ++h) valid[h] = fq_load_point(x, base0 + 8 * h + g, B, bound, x01[h]);
        fq_group_issue<0, SAFE>(table, s.lv[t], x01, align_corners, raw, frac);
    }
#endif
    for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        const int64_t base = tile * VG_TILE + warp * 16;
#if !VG_PIPE
        if (base >= B) continue;
#pragma unroll
        for (int h = 0; h < 2; ++h) valid[h] = fq_load_point(x, base + 8 * h + g, B, bound, x01[h]);
        fq_group_issue<0, SAFE>(table, s.lv[t], x01, align_corners, raw, frac);
#endif

        // ================================================================ forward (as field.cu) + corner cache
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
                if (kt >= VG_G0) {
#pragma unroll
                    for (int p = 0; p < 4; ++p)
                        cache[p][g + 8 * h][4 * (kt - VG_G0) + t] = make_float4(v[2 * p].x, v[2 * p].y, v[2 * p + 1].x, v[2 * p + 1].y);
                }
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
            for (int nt = 0; nt < 8; ++nt) fq_mma3(acc[nt], ahi, alo, s.wF1[(nt * 4 + kt) * 32 + lane]);
        }
#if !VG_PIPE
        // the backward's re-gather of the four coarse levels goes out now and lands during layer 2 / the dFeat products
        fq_group_issue<0, SAFE>(table, s.lv[t], x01, align_corners, raw, frac);
#endif

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
            for (int nt = 0; nt < 2; ++nt) fq_mma3(out[nt], ahi, alo, s.wF2[(nt * 8 + kt) * 32 + lane]);
        }
        // sigma of rows g, g+8 lives on quad lane 0: share it with the quad (the backward needs it on every lane)
        float sg[2], go0[2];
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const int64_t b = base + 8 * h + g;
            sg[h] = __shfl_sync(0xffffffffu, expf(out[0][2 * h]), lane & ~3);   // trunc_exp forward, activation.py:9-11
            const bool inb = b < B;
            if (inb && t == 0) sigma[b] = sg[h];
            if (inb && geo) {
                float *gp = geo + b * (FQ_OUT - 1);
#pragma unroll
                for (int nt = 0; nt < 2; ++nt) {
                    const int col = 8 * nt + 2 * t;
                    if (col > 0) gp[col - 1] = out[nt][2 * h];
                    gp[col] = out[nt][2 * h + 1];
                }
            }
            // upstream factor u = a + b sigma, times trunc_exp' = exp(clamp(o0, -15, 15)) = clamp(sigma)  (activation.py:13-17)
            const float ua = (inb && coef_a) ? __ldg(coef_a + b) : a0, ub = (inb && coef_b) ? __ldg(coef_b + b) : b0;
            go0[h] = inb ? fmaf(ub, sg[h], ua) * fminf(fmaxf(sg[h], e_lo), e_hi) : 0.0f;
        }

        // ================================================================ backward to x
        float dhm[8][4];
#pragma unroll
        for (int nt = 0; nt < 8; ++nt) {
            const float2 w = *reinterpret_cast<const float2 *>(&s.w1row0[8 * nt + 2 * t]);
            const uint32_t m0 = mbits[0] >> (2 * nt), m1 = mbits[1] >> (2 * nt);
            dhm[nt][0] = (m0 & 1u) ? go0[0] * w.x : 0.0f;
            dhm[nt][1] = (m0 & 2u) ? go0[0] * w.y : 0.0f;
            dhm[nt][2] = (m1 & 1u) ? go0[1] * w.x : 0.0f;
            dhm[nt][3] = (m1 & 2u) ? go0[1] * w.y : 0.0f;
        }
        float gx[2][3];
#pragma unroll
        for (int h = 0; h < 2; ++h) gx[h][0] = gx[h][1] = gx[h][2] = 0.0f;
        __syncwarp();   // the warp's cache rows were written by the lanes that read them, but keep the order explicit
#if VG_PIPE
        // the NEXT tile's points and first level group go out now: its gathers fly during this tile's backward
        float xn[2][3];
        bool vn[2];
        {
            const int64_t nbase = (tile + gridDim.x) * VG_TILE + warp * 16;
#pragma unroll
            for (int h = 0; h < 2; ++h) vn[h] = fq_load_point(x, nbase + 8 * h + g, B, bound, xn[h]);
            fq_group_issue<0, SAFE>(table, s.lv[t], xn, align_corners, raw, frac);
        }
#endif
        // cached level groups
#pragma unroll 1
        for (int nt = VG_G0; nt < 4; ++nt) {
            float df[4];
            vg_dfeat(s.wB1, nt, lane, dhm, df);
            const float scale = s.lv[4 * nt + t].scale;
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                uint32_t cell[3];
                float fr[3];
                fq_cell(x01[h], scale, align_corners, cell, fr);
                float2 v[8];
#pragma unroll
                for (int p = 0; p < 4; ++p) {
                    const float4 c = cache[p][g + 8 * h][4 * (nt - VG_G0) + t];
                    v[2 * p] = make_float2(c.x, c.y);
                    v[2 * p + 1] = make_float2(c.z, c.w);
                }
                vg_contract(v, fr, scale, valid[h] ? df[2 * h] : 0.0f, valid[h] ? df[2 * h + 1] : 0.0f, gx[h]);
            }
        }
#if !VG_PIPE
        // coarse level group 0 (re-gathered above)
        {
            float df[4];
            vg_dfeat(s.wB1, 0, lane, dhm, df);
            const float scale = s.lv[t].scale;
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                float2 v[8];
                fq_select(raw[h], v);
                vg_contract(v, frac[h], scale, valid[h] ? df[2 * h] : 0.0f, valid[h] ? df[2 * h + 1] : 0.0f, gx[h]);
            }
        }
#endif
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
        __syncwarp();
#if VG_PIPE
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            valid[h] = vn[h];
            x01[h][0] = xn[h][0]; x01[h][1] = xn[h][1]; x01[h][2] = xn[h][2];
        }
#endif
    }
}

__global__ void __launch_bounds__(VG_THREADS, 1)
k_field_value_grad(const float *__restrict__ x, float bound, const float *__restrict__ table,
                   const int32_t *__restrict__ offsets, float S, uint32_t H, uint32_t gridtype, int align_corners,
                   const float *__restrict__ W0, const float *__restrict__ W1, const float *__restrict__ coef_a,
                   const float *__restrict__ coef_b, float a0, float b0, float *__restrict__ sigma, float *__restrict__ geo,
                   float *__restrict__ grad_x, int64_t B) {
    extern __shared__ __align__(16) unsigned char vg_smem_raw[];
    FieldVGSmem &s = *reinterpret_cast<FieldVGSmem *>(vg_smem_raw);
    const int tid = threadIdx.x;
    fq_stage_levels(s.lv, &s.all_safe, offsets, S, H, gridtype, align_corners);
    for (int i = tid; i < 1024; i += VG_THREADS) {  // forward layer 1 (field.cu k_field_forward)
        const int ln = i & 31, kt = (i >> 5) & 3, nt = i >> 7, gg = ln >> 2, tt = ln & 3;
        const float *w = W0 + (8 * nt + gg) * FQ_F + 8 * kt + 2 * tt;
        s.wF1[i] = fq_split_pair(w[0], w[1]);
    }
    for (int i = tid; i < 512; i += VG_THREADS) {   // forward layer 2
        const int ln = i & 31, kt = (i >> 5) & 7, nt = i >> 8, gg = ln >> 2, tt = ln & 3;
        const float *w = W1 + (8 * nt + gg) * FQ_HID + 8 * kt + 2 * tt;
        s.wF2[i] = fq_split_pair(w[0], w[1]);
    }
    for (int i = tid; i < 1024; i += VG_THREADS) {  // backward dFeat = dH W0 (field.cu k_field_backward)
        const int ln = i & 31, kt = (i >> 5) & 7, nt = i >> 8, gg = ln >> 2, tt = ln & 3;
        const int n = 8 * nt + gg;
        s.wB1[i] = fq_split_pair(W0[(8 * kt + 2 * tt) * FQ_F + n], W0[(8 * kt + 2 * tt + 1) * FQ_F + n]);
    }
    if (tid < FQ_HID) s.w1row0[tid] = W1[tid];
    __syncthreads();
    if (s.all_safe) vg_tiles<true>(s, x, bound, table, align_corners, coef_a, coef_b, a0, b0, sigma, geo, grad_x, B);
    else vg_tiles<false>(s, x, bound, table, align_corners, coef_a, coef_b, a0, b0, sigma, geo, grad_x, B);
}

// default implementation: the single-table-pass tcgen05 kernel of dfield5.cu; ngp_set_field_impl(0) / NGP_B200_FIELD_IMPL=mma
// selects the mma.sync kernel above
int ngp_field_impl_tc5();
int d5_field_query_value_grad(const float *x, float bound, const float *embeddings, const int32_t *offsets, float S, uint32_t H,
                              uint32_t gridtype, int align_corners, const float *W0, const float *W1, const float *coef_a,
                              const float *coef_b, float a0, float b0, float *sigma, float *geo, float *grad_x,
                              uint64_t *relu_mask, int64_t B, cudaStream_t stream);

extern "C" int ngp_field_query_value_grad(const float *x, float bound, const float *embeddings, const int32_t *offsets,
                                          float S, uint32_t H, uint32_t gridtype, int align_corners, const float *W0,
                                          const float *W1, const float *coef_a, const float *coef_b, float a0, float b0,
                                          float *sigma, float *geo, float *grad_x, int64_t B, ngp_stream_t stream) {
    NGP_REQUIRE_DEVICE(x, "ngp_field_query_value_grad");
    if (B == 0) return NGP_OK;
    NGP_REQUIRE(x && embeddings && offsets && W0 && W1 && sigma && grad_x, NGP_ERR_ARG,
                "ngp_field_query_value_grad: null pointer");
    NGP_REQUIRE(bound > 0.0f && B >= 0, NGP_ERR_ARG, "ngp_field_query_value_grad: bad bound / B");
    NGP_REQUIRE(((uintptr_t)embeddings & 15u) == 0, NGP_ERR_ARG, "ngp_field_query_value_grad: embeddings must be 16-byte aligned");
    if (ngp_field_impl_tc5())
        return d5_field_query_value_grad(x, bound, embeddings, offsets, S, H, gridtype, align_corners, W0, W1, coef_a, coef_b, a0,
                                         b0, sigma, geo, grad_x, nullptr, B, (cudaStream_t)stream);
    const int smem = (int)sizeof(FieldVGSmem);
    if (cudaFuncSetAttribute(k_field_value_grad, cudaFuncAttributeMaxDynamicSharedMemorySize, smem) != cudaSuccess) {
        ngp_set_error("ngp_field_query_value_grad: cannot reserve %d B of shared memory", smem);
        return NGP_ERR_CUDA;
    }
    int dev = 0, sms = NGP_SM_COUNT;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const int64_t tiles = ngp_div_up(B, VG_TILE);
    const int grid = (int)(tiles < sms ? tiles : sms);
    k_field_value_grad<<<grid, VG_THREADS, smem, (cudaStream_t)stream>>>(x, bound, embeddings, offsets, S, H, gridtype,
                                                                         align_corners, W0, W1, coef_a, coef_b, a0, b0, sigma,
                                                                         geo, grad_x, B);
    NGP_CHECK_LAUNCH("ngp_field_query_value_grad");
    return NGP_OK;
}

extern "C" int ngp_field_query_forward(const float *x, float bound, const float *embeddings, const int32_t *offsets, float S,
                                       uint32_t H, uint32_t gridtype, int align_corners, const float *W0, const float *W1,
                                       float *sigma, float *geo, uint64_t *relu_mask, int64_t B, ngp_stream_t stream);

// ---- density forward that also returns the Jacobian rows: what autograd's density(x) runs when x requires a gradient.
// The backward of sigma = f(x_i) is then grad_sigma_i * jac_i: one streaming kernel instead of a second table pass.
// `out` may be `jac` itself (in place): every element is read and written by the same thread, no read-only path for it
__global__ void k_scale_rows3(const float *jac, const float *__restrict__ gs, float *out, int64_t n3) {
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n3; i += (int64_t)gridDim.x * blockDim.x)
        out[i] = __ldg(gs + i / 3) * jac[i];
}

extern "C" int ngp_field_query_forward_jacobian(const float *x, float bound, const float *embeddings, const int32_t *offsets,
                                                float S, uint32_t H, uint32_t gridtype, int align_corners, const float *W0,
                                                const float *W1, float *sigma, float *geo, uint64_t *relu_mask, float *jac,
                                                int64_t B, ngp_stream_t stream) {
    NGP_REQUIRE_DEVICE(x, "ngp_field_query_forward_jacobian");
    if (B == 0) return NGP_OK;
    NGP_REQUIRE(x && embeddings && offsets && W0 && W1 && sigma && jac, NGP_ERR_ARG, "ngp_field_query_forward_jacobian: null pointer");
    NGP_REQUIRE(bound > 0.0f && B >= 0, NGP_ERR_ARG, "ngp_field_query_forward_jacobian: bad bound / B");
    NGP_REQUIRE(((uintptr_t)embeddings & 15u) == 0, NGP_ERR_ARG, "ngp_field_query_forward_jacobian: embeddings must be 16-byte aligned");
    if (ngp_field_impl_tc5())
        return d5_field_query_value_grad(x, bound, embeddings, offsets, S, H, gridtype, align_corners, W0, W1, nullptr, nullptr, 1.0f,
                                         0.0f, sigma, geo, jac, relu_mask, B, (cudaStream_t)stream);
    // mma.sync cross-check path: the forward kernel (sigma, geo, mask) and the single-pass kernel for the Jacobian rows
    int rc = ngp_field_query_forward(x, bound, embeddings, offsets, S, H, gridtype, align_corners, W0, W1, sigma, geo, relu_mask, B,
                                     stream);
    if (rc != NGP_OK) return rc;
    return ngp_field_query_value_grad(x, bound, embeddings, offsets, S, H, gridtype, align_corners, W0, W1, nullptr, nullptr, 1.0f,
                                      0.0f, sigma, nullptr, jac, B, stream);
}

extern "C" int ngp_field_query_backward_jacobian(const float *jac, const float *grad_sigma, float *grad_x, int64_t B,
                                                 ngp_stream_t stream) {
    NGP_REQUIRE_DEVICE(jac, "ngp_field_query_backward_jacobian");
    if (B == 0) return NGP_OK;
    NGP_REQUIRE(jac && grad_sigma && grad_x && B >= 0, NGP_ERR_ARG, "ngp_field_query_backward_jacobian: null pointer / bad B");
    const int64_t n3 = 3 * B;
    const int64_t blocks = ngp_div_up(n3, 256);
    k_scale_rows3<<<(int)(blocks < 148 * 16 ? blocks : 148 * 16), 256, 0, (cudaStream_t)stream>>>(jac, grad_sigma, grad_x, n3);
    NGP_CHECK_LAUNCH("ngp_field_query_backward_jacobian");
    return NGP_OK;
}
