// renderfield.cu — kernels + C ABI of the fused per-sample field (building blocks in renderfield.cuh).
#include "renderfield.cuh"

// ------------------------------------------------------------------------------------------------ forward
// masks (optional, consumed by the backward): u16 [M][3][4]: (sigma hidden | colour hidden 1 | colour hidden 2) x quad
// lane t; bit (2*nt + c) of entry t = unit 8*nt + 2*t + c.
template <bool SAFE>
__device__ __forceinline__ void rf_forward_tiles(const RenderFieldSmem &s, const float *__restrict__ x,
                                                 const float *__restrict__ dirs, float bound,
                                                 const float *__restrict__ table, int align_corners,
                                                 float *__restrict__ sigma, float *__restrict__ rgb,
                                                 uint16_t *__restrict__ masks, int64_t B) {
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, g = lane >> 2, t = lane & 3;
    const int64_t ntiles = (B + RF_TILE - 1) / RF_TILE;
    for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        const int64_t base = tile * RF_TILE + warp * 16;
        if (base >= B) continue;
        float x01[2][3];
        bool valid[2];
#pragma unroll
        for (int h = 0; h < 2; ++h) valid[h] = fq_load_point(x, base + 8 * h + g, B, bound, x01[h]);
        float dir[2][3];
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const int64_t b = base + 8 * h + g;
#pragma unroll
            for (int a = 0; a < 3; ++a) dir[h][a] = b < B ? __ldg(dirs + b * 3 + a) : 0.0f;
        }
        float o[2][4], o3[1][4];
        uint32_t m_sig[2], m_c1[2], m_c2[2];
        rf_eval16<SAFE>(s, table, align_corners, lane, x01, valid, dir, o, o3, m_sig, m_c1, m_c2);

        // ---- outputs
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const int64_t b = base + 8 * h + g;
            if (b >= B) continue;
            if (masks) {
                masks[(b * 3 + 0) * 4 + t] = (uint16_t)m_sig[h];
                masks[(b * 3 + 1) * 4 + t] = (uint16_t)m_c1[h];
                masks[(b * 3 + 2) * 4 + t] = (uint16_t)m_c2[h];
            }
            if (t == 0) sigma[b] = expf(o[0][2 * h]);  // trunc_exp forward (activation.py:9-11)
            const float v0 = 1.0f / (1.0f + expf(-o3[0][2 * h])), v1 = 1.0f / (1.0f + expf(-o3[0][2 * h + 1]));
            if (t == 0) {
                rgb[b * 3] = v0;
                rgb[b * 3 + 1] = v1;
            } else if (t == 1) {
                rgb[b * 3 + 2] = v0;
            }
        }
    }
}

__global__ void __launch_bounds__(RF_THREADS, 1)
k_render_field_forward(const float *__restrict__ x, const float *__restrict__ dirs, float bound,
                       const float *__restrict__ table, const int32_t *__restrict__ offsets, float S, uint32_t H,
                       uint32_t gridtype, int align_corners, const float *__restrict__ Ws0,
                       const float *__restrict__ Ws1, const float *__restrict__ Wc0, const float *__restrict__ Wc1,
                       const float *__restrict__ Wc2, float *__restrict__ sigma, float *__restrict__ rgb,
                       uint16_t *__restrict__ masks, int64_t B) {
    extern __shared__ __align__(16) unsigned char rf_smem_raw[];
    RenderFieldSmem &s = *reinterpret_cast<RenderFieldSmem *>(rf_smem_raw);
    const int tid = threadIdx.x;
    fq_stage_levels(s.lv, &s.all_safe, offsets, S, H, gridtype, align_corners);
    rf_stage_forward_weights(s, Ws0, Ws1, Wc0, Wc1, Wc2);
    __syncthreads();
    if (s.all_safe) rf_forward_tiles<true>(s, x, dirs, bound, table, align_corners, sigma, rgb, masks, B);
    else rf_forward_tiles<false>(s, x, dirs, bound, table, align_corners, sigma, rgb, masks, B);
}

// ------------------------------------------------------------------------------------------------ backward to (x, d)
// out[nt] = (in .* mask) * W for a 64-wide accumulator-fragment input; W fragments [nt NT][kt 8][lane]
template <int NT>
__device__ __forceinline__ void rf_masked_layer(const float (&in)[8][4], const uint32_t (&mbits)[2],
                                                const RfFrag *__restrict__ frags, int lane, float (&out)[NT][4]) {
#pragma unroll
    for (int nt = 0; nt < NT; ++nt)
#pragma unroll
        for (int c = 0; c < 4; ++c) out[nt][c] = 0.0f;
#pragma unroll
    for (int kt = 0; kt < 8; ++kt) {
        const uint32_t m0 = mbits[0] >> (2 * kt), m1 = mbits[1] >> (2 * kt);
        uint32_t ahi[4], alo[4];
        fq_split((m0 & 1u) ? in[kt][0] : 0.0f, ahi[0], alo[0]);
        fq_split((m1 & 1u) ? in[kt][2] : 0.0f, ahi[1], alo[1]);
        fq_split((m0 & 2u) ? in[kt][1] : 0.0f, ahi[2], alo[2]);
        fq_split((m1 & 2u) ? in[kt][3] : 0.0f, ahi[3], alo[3]);
#pragma unroll
        for (int nt = 0; nt < NT; ++nt) fq_mma3(out[nt], ahi, alo, frags[(nt * 8 + kt) * 32 + lane]);
    }
}

template <bool SAFE>
__device__ __forceinline__ void rf_backward_tiles(const RenderFieldSmem &s, const float *__restrict__ x,
                                                  const float *__restrict__ dirs, float bound,
                                                  const float *__restrict__ table, int align_corners,
                                                  const float *__restrict__ sigma, const float *__restrict__ rgb,
                                                  const uint16_t *__restrict__ masks,
                                                  const float *__restrict__ grad_sigma,
                                                  const float *__restrict__ grad_rgb, float *__restrict__ grad_x,
                                                  float *__restrict__ grad_d, int64_t B) {
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, g = lane >> 2, t = lane & 3;
    const float e_lo = expf(-15.0f), e_hi = expf(15.0f);
    const float den = 2.0f * bound;
    const int64_t ntiles = (B + RF_TILE - 1) / RF_TILE;
    for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        const int64_t base = tile * RF_TILE + warp * 16;
        if (base >= B) continue;
        float x01[2][3], dir[2][3], go0[2], d3[2];  // d3: dL/d(colour logit t) of the lane's two rows (t < 3)
        bool valid[2];
        uint32_t m_sig[2], m_c1[2], m_c2[2];
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const int64_t b = base + 8 * h + g;
            valid[h] = fq_load_point(x, b, B, bound, x01[h]);
            const bool inb = b < B;
            go0[h] = inb ? __ldg(grad_sigma + b) * fminf(fmaxf(__ldg(sigma + b), e_lo), e_hi) : 0.0f;
            float c = 0.0f, gc = 0.0f;
            if (inb && t < 3) {
                c = __ldg(rgb + b * 3 + t);
                gc = __ldg(grad_rgb + b * 3 + t);
            }
            d3[h] = gc * c * (1.0f - c);  // sigmoid'
#pragma unroll
            for (int a = 0; a < 3; ++a) dir[h][a] = inb ? __ldg(dirs + b * 3 + a) : 0.0f;
            m_sig[h] = inb ? (uint32_t)__ldg(masks + (b * 3 + 0) * 4 + t) : 0u;
            m_c1[h] = inb ? (uint32_t)__ldg(masks + (b * 3 + 1) * 4 + t) : 0u;
            m_c2[h] = inb ? (uint32_t)__ldg(masks + (b * 3 + 2) * 4 + t) : 0u;
        }
        // ---- colour layer 3 backward: dHc2 = dRGB W_c2 (K = 3 padded to one k-tile)
        float dh2[8][4];
        {
            uint32_t ahi[4], alo[4];
            fq_split(d3[0], ahi[0], alo[0]);
            fq_split(d3[1], ahi[1], alo[1]);
            ahi[2] = ahi[3] = alo[2] = alo[3] = 0u;
#pragma unroll
            for (int nt = 0; nt < 8; ++nt) {
                dh2[nt][0] = dh2[nt][1] = dh2[nt][2] = dh2[nt][3] = 0.0f;
                fq_mma3(dh2[nt], ahi, alo, s.c2[nt * 32 + lane]);
            }
        }
        // ---- colour layers 2, 1 backward
        float dh1[8][4], din[4][4];
        rf_masked_layer<8>(dh2, m_c2, s.c1, lane, dh1);
        rf_masked_layer<4>(dh1, m_c1, s.c0, lane, din);  // din[0..1] = dL/dSH 8nt+2t(+1); din[2..3] = dL/do 8(nt-2)+2t(+1)

        // ---- SH backward: the lane holds dL/dSH at indices {2t, 2t+1, 8+2t, 9+2t}; J^T v, summed over the quad below
        float gdir[2][3];
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            float jx[16], jy[16], jz[16];
            ngp_sh4_jac(dir[h][0], dir[h][1], dir[h][2], jx, jy, jz);
            gdir[h][0] = gdir[h][1] = gdir[h][2] = 0.0f;
#pragma unroll
            for (int nt = 0; nt < 2; ++nt)
#pragma unroll
                for (int c = 0; c < 2; ++c) {
                    const int o8 = 8 * nt + c;
                    const float v = din[nt][2 * h + c];
                    gdir[h][0] = fmaf(v, rf_sel4(t, jx[o8], jx[o8 + 2], jx[o8 + 4], jx[o8 + 6]), gdir[h][0]);
                    gdir[h][1] = fmaf(v, rf_sel4(t, jy[o8], jy[o8 + 2], jy[o8 + 4], jy[o8 + 6]), gdir[h][1]);
                    gdir[h][2] = fmaf(v, rf_sel4(t, jz[o8], jz[o8 + 2], jz[o8 + 4], jz[o8 + 6]), gdir[h][2]);
                }
        }
        // ---- sigma net backward: dO = [dL/dsigma * exp'(.) | dL/dgeo], dHs = dO W_s1 (K = 16), ReLU mask
        if (t == 0) {
            din[2][0] += go0[0];  // column 0 (its colour weight is zero, so din holds 0 there)
            din[2][2] += go0[1];
        }
        float dhm[8][4];
#pragma unroll
        for (int nt = 0; nt < 8; ++nt) dhm[nt][0] = dhm[nt][1] = dhm[nt][2] = dhm[nt][3] = 0.0f;
#pragma unroll
        for (int kt = 0; kt < 2; ++kt) {
            uint32_t ahi[4], alo[4];
            fq_split(din[2 + kt][0], ahi[0], alo[0]);
            fq_split(din[2 + kt][2], ahi[1], alo[1]);
            fq_split(din[2 + kt][1], ahi[2], alo[2]);
            fq_split(din[2 + kt][3], ahi[3], alo[3]);
#pragma unroll
            for (int nt = 0; nt < 8; ++nt) fq_mma3(dhm[nt], ahi, alo, s.s1[(nt * 2 + kt) * 32 + lane]);
        }
#pragma unroll
        for (int nt = 0; nt < 8; ++nt) {
            const uint32_t m0 = m_sig[0] >> (2 * nt), m1 = m_sig[1] >> (2 * nt);
            dhm[nt][0] = (m0 & 1u) ? dhm[nt][0] : 0.0f;
            dhm[nt][1] = (m0 & 2u) ? dhm[nt][1] : 0.0f;
            dhm[nt][2] = (m1 & 1u) ? dhm[nt][2] : 0.0f;
            dhm[nt][3] = (m1 & 2u) ? dhm[nt][3] : 0.0f;
        }
        // ---- encoder input-backward, one level group at a time (fieldcore.cuh)
        float gx[2][3];
#pragma unroll
        for (int h = 0; h < 2; ++h) gx[h][0] = gx[h][1] = gx[h][2] = 0.0f;
        fq_backward_group<0, SAFE>(s.lv, s.s0, table, 0, lane, x01, valid, align_corners, dhm, gx);
#pragma unroll 1
        for (int nt = 1; nt < 4; ++nt) fq_backward_group<1, SAFE>(s.lv, s.s0, table, nt, lane, x01, valid, align_corners, dhm, gx);

#pragma unroll
        for (int h = 0; h < 2; ++h) {
#pragma unroll
            for (int a = 0; a < 3; ++a) {
                gx[h][a] += __shfl_xor_sync(0xffffffffu, gx[h][a], 1);
                gx[h][a] += __shfl_xor_sync(0xffffffffu, gx[h][a], 2);
                gdir[h][a] += __shfl_xor_sync(0xffffffffu, gdir[h][a], 1);
                gdir[h][a] += __shfl_xor_sync(0xffffffffu, gdir[h][a], 2);
            }
            const int64_t b = base + 8 * h + g;
            if (b < B && t < 3) {
                grad_x[b * 3 + t] = __fdiv_rn(t == 0 ? gx[h][0] : (t == 1 ? gx[h][1] : gx[h][2]), den);
                if (grad_d) grad_d[b * 3 + t] = t == 0 ? gdir[h][0] : (t == 1 ? gdir[h][1] : gdir[h][2]);
            }
        }
    }
}

__global__ void __launch_bounds__(RF_THREADS, 1)
k_render_field_backward(const float *__restrict__ x, const float *__restrict__ dirs, float bound,
                        const float *__restrict__ table, const int32_t *__restrict__ offsets, float S, uint32_t H,
                        uint32_t gridtype, int align_corners, const float *__restrict__ Ws0,
                        const float *__restrict__ Ws1, const float *__restrict__ Wc0, const float *__restrict__ Wc1,
                        const float *__restrict__ Wc2, const float *__restrict__ sigma, const float *__restrict__ rgb,
                        const uint16_t *__restrict__ masks, const float *__restrict__ grad_sigma,
                        const float *__restrict__ grad_rgb, float *__restrict__ grad_x, float *__restrict__ grad_d,
                        int64_t B) {
    extern __shared__ __align__(16) unsigned char rf_smem_raw[];
    RenderFieldSmem &s = *reinterpret_cast<RenderFieldSmem *>(rf_smem_raw);
    const int tid = threadIdx.x;
    fq_stage_levels(s.lv, &s.all_safe, offsets, S, H, gridtype, align_corners);
    // every product below is (row fragment) x W: B[k][n] = W[k-th unit of the upstream layer][n-th unit downstream],
    // k slots (t, t+4) of k-tile kt = upstream units 8kt+2t, 8kt+2t+1 (the accumulator-fragment columns)
    for (int i = tid; i < 1024; i += RF_THREADS) {  // dFeat = dHs W_s0: [nt 4][kt 8]
        const int ln = i & 31, kt = (i >> 5) & 7, nt = i >> 8, gg = ln >> 2, tt = ln & 3, n = 8 * nt + gg;
        fq_store_frag(s.s0[i], Ws0[(8 * kt + 2 * tt) * FQ_F + n], Ws0[(8 * kt + 2 * tt + 1) * FQ_F + n]);
    }
    for (int i = tid; i < 512; i += RF_THREADS) {  // dHs = dO W_s1: [nt 8][kt 2], k = output columns
        const int ln = i & 31, kt = (i >> 5) & 1, nt = i >> 6, gg = ln >> 2, tt = ln & 3, n = 8 * nt + gg;
        fq_store_frag(s.s1[i], Ws1[(8 * kt + 2 * tt) * FQ_HID + n], Ws1[(8 * kt + 2 * tt + 1) * FQ_HID + n]);
    }
    for (int i = tid; i < 1024; i += RF_THREADS) {  // dIn = dHc1 W_c0: [nt 4][kt 8]; n-tiles 0,1 = SH, 2,3 = o columns
        const int ln = i & 31, kt = (i >> 5) & 7, nt = i >> 8, gg = ln >> 2, tt = ln & 3;
        const float *w0 = Wc0 + (8 * kt + 2 * tt) * RF_CIN, *w1 = w0 + RF_CIN;
        float b0, b1;
        if (nt < 2) {
            b0 = w0[8 * nt + gg];
            b1 = w1[8 * nt + gg];
        } else {
            const int col = 8 * (nt - 2) + gg;
            b0 = col == 0 ? 0.0f : w0[16 + col - 1];
            b1 = col == 0 ? 0.0f : w1[16 + col - 1];
        }
        fq_store_frag(s.c0[i], b0, b1);
    }
    for (int i = tid; i < 2048; i += RF_THREADS) {  // dHc1 = dHc2 W_c1: [nt 8][kt 8]
        const int ln = i & 31, kt = (i >> 5) & 7, nt = i >> 8, gg = ln >> 2, tt = ln & 3, n = 8 * nt + gg;
        fq_store_frag(s.c1[i], Wc1[(8 * kt + 2 * tt) * FQ_HID + n], Wc1[(8 * kt + 2 * tt + 1) * FQ_HID + n]);
    }
    for (int i = tid; i < 256; i += RF_THREADS) {  // dHc2 = dRGB W_c2: [nt 8], k slot t = colour channel t (< 3)
        const int ln = i & 31, nt = i >> 5, gg = ln >> 2, tt = ln & 3;
        fq_store_frag(s.c2[i], tt < 3 ? Wc2[tt * FQ_HID + 8 * nt + gg] : 0.0f, 0.0f);
    }
    __syncthreads();
    if (s.all_safe)
        rf_backward_tiles<true>(s, x, dirs, bound, table, align_corners, sigma, rgb, masks, grad_sigma, grad_rgb, grad_x, grad_d, B);
    else
        rf_backward_tiles<false>(s, x, dirs, bound, table, align_corners, sigma, rgb, masks, grad_sigma, grad_rgb, grad_x, grad_d, B);
}

// ------------------------------------------------------------------------------------------------ C ABI
// default implementation: the warp-specialised tcgen05 kernels of cfield5.cu; NGP_B200_FIELD_IMPL=mma selects the
// mma.sync kernels above (kept as the independent fp32 cross-check)
int ngp_field_impl_tc5();
int c5_field_color_forward(const float *xyzs, const float *dirs, float bound, const float *embeddings,
                           const int32_t *offsets, float S, uint32_t H, uint32_t gridtype, int align_corners,
                           const float *Ws0, const float *Ws1, const float *Wc0, const float *Wc1, const float *Wc2,
                           float *sigma, float *rgb, void *masks, int64_t M, cudaStream_t stream);
int c5_field_color_backward(const float *xyzs, const float *dirs, float bound, const float *embeddings,
                            const int32_t *offsets, float S, uint32_t H, uint32_t gridtype, int align_corners,
                            const float *Ws0, const float *Ws1, const float *Wc0, const float *Wc1, const float *Wc2,
                            const float *sigma, const float *rgb, const void *masks, const float *grad_sigma,
                            const float *grad_rgb, float *grad_xyzs, float *grad_dirs, int64_t M, cudaStream_t stream);

template <typename K>
static int rf_launch_cfg(K kernel, int64_t B, int *grid) {
    const int smem = (int)sizeof(RenderFieldSmem);
    if (cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem) != cudaSuccess) {
        ngp_set_error("render field: cannot reserve %d B of shared memory", smem);
        return NGP_ERR_CUDA;
    }
    int dev = 0, sms = NGP_SM_COUNT;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const int64_t tiles = ngp_div_up(B, RF_TILE);
    *grid = (int)(tiles < sms ? tiles : sms);
    return NGP_OK;
}

extern "C" int ngp_field_color_forward(const float *xyzs, const float *dirs, float bound, const float *embeddings,
                                       const int32_t *offsets, float S, uint32_t H, uint32_t gridtype,
                                       int align_corners, const float *Ws0, const float *Ws1, const float *Wc0,
                                       const float *Wc1, const float *Wc2, float *sigma, float *rgb, void *masks,
                                       int64_t M, ngp_stream_t stream) {
    NGP_REQUIRE_DEVICE(xyzs, "ngp_field_color_forward");
    if (M == 0) return NGP_OK;
    NGP_REQUIRE(xyzs && dirs && embeddings && offsets && Ws0 && Ws1 && Wc0 && Wc1 && Wc2 && sigma && rgb, NGP_ERR_ARG,
                "ngp_field_color_forward: null pointer");
    NGP_REQUIRE(bound > 0.0f && M >= 0, NGP_ERR_ARG, "ngp_field_color_forward: bad bound / M");
    NGP_REQUIRE(((uintptr_t)embeddings & 15u) == 0, NGP_ERR_ARG, "ngp_field_color_forward: embeddings must be 16-byte aligned");
    if (ngp_field_impl_tc5())
        return c5_field_color_forward(xyzs, dirs, bound, embeddings, offsets, S, H, gridtype, align_corners, Ws0, Ws1, Wc0,
                                      Wc1, Wc2, sigma, rgb, masks, M, (cudaStream_t)stream);
    int grid = 1;
    int rc = rf_launch_cfg(k_render_field_forward, M, &grid);
    if (rc) return rc;
    k_render_field_forward<<<grid, RF_THREADS, sizeof(RenderFieldSmem), (cudaStream_t)stream>>>(
        xyzs, dirs, bound, embeddings, offsets, S, H, gridtype, align_corners, Ws0, Ws1, Wc0, Wc1, Wc2, sigma, rgb,
        (uint16_t *)masks, M);
    NGP_CHECK_LAUNCH("ngp_field_color_forward");
    return NGP_OK;
}

extern "C" int ngp_field_color_backward(const float *xyzs, const float *dirs, float bound, const float *embeddings,
                                        const int32_t *offsets, float S, uint32_t H, uint32_t gridtype,
                                        int align_corners, const float *Ws0, const float *Ws1, const float *Wc0,
                                        const float *Wc1, const float *Wc2, const float *sigma, const float *rgb,
                                        const void *masks, const float *grad_sigma, const float *grad_rgb,
                                        float *grad_xyzs, float *grad_dirs, int64_t M, ngp_stream_t stream) {
    NGP_REQUIRE_DEVICE(xyzs, "ngp_field_color_backward");
    if (M == 0) return NGP_OK;
    NGP_REQUIRE(xyzs && dirs && embeddings && offsets && Ws0 && Ws1 && Wc0 && Wc1 && Wc2 && sigma && rgb && masks &&
                    grad_sigma && grad_rgb && grad_xyzs,
                NGP_ERR_ARG, "ngp_field_color_backward: null pointer");
    NGP_REQUIRE(bound > 0.0f && M >= 0, NGP_ERR_ARG, "ngp_field_color_backward: bad bound / M");
    NGP_REQUIRE(((uintptr_t)embeddings & 15u) == 0, NGP_ERR_ARG, "ngp_field_color_backward: embeddings must be 16-byte aligned");
    if (ngp_field_impl_tc5())
        return c5_field_color_backward(xyzs, dirs, bound, embeddings, offsets, S, H, gridtype, align_corners, Ws0, Ws1, Wc0,
                                       Wc1, Wc2, sigma, rgb, masks, grad_sigma, grad_rgb, grad_xyzs, grad_dirs, M,
                                       (cudaStream_t)stream);
    int grid = 1;
    int rc = rf_launch_cfg(k_render_field_backward, M, &grid);
    if (rc) return rc;
    k_render_field_backward<<<grid, RF_THREADS, sizeof(RenderFieldSmem), (cudaStream_t)stream>>>(
        xyzs, dirs, bound, embeddings, offsets, S, H, gridtype, align_corners, Ws0, Ws1, Wc0, Wc1, Wc2, sigma, rgb,
        (const uint16_t *)masks, grad_sigma, grad_rgb, grad_xyzs, grad_dirs, M);
    NGP_CHECK_LAUNCH("ngp_field_color_backward");
    return NGP_OK;
}
