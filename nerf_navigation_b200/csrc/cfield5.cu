// cfield5.cu — warp-specialised tcgen05 kernels of the fused per-sample field (building blocks in cfield5.cuh):
// forward (x, d) -> (sigma, rgb) and backward to (x, d) for the ray paths of NeRFRenderer.run_cuda (renderer.py:303-305,
// 352-354) / the pose estimator.  These are the default implementation behind ngp_field_color_forward/backward
// (renderfield.cu keeps the mma.sync 3xTF32 kernels as the cross-check, NGP_B200_FIELD_IMPL=mma).
#include <stdlib.h>
#include <string.h>

#include "cfield5.cuh"

// 1: tcgen05 kernels (default), 0: legacy mma.sync kernels
static int g_field_impl = -1;
int ngp_field_impl_tc5() {
    if (g_field_impl < 0) {
        const char *e = getenv("NGP_B200_FIELD_IMPL");
        g_field_impl = (e && strcmp(e, "mma") == 0) ? 0 : 1;
    }
    return g_field_impl;
}
extern "C" int ngp_set_field_impl(int impl) {
    const int prev = ngp_field_impl_tc5();
    if (impl >= 0) g_field_impl = impl ? 1 : 0;
    return prev;
}

// -DC5_STATS (dev builds only, scripts/c5_roles.py): cycles each role of the forward spends inside its barrier waits and
// in total, summed over the CTAs' reporting threads (one lane per warp).  [0,1] gather: wait on f_empty, total;
// [2,3] chain: wait on d_full, total; [4..7] MMA thread: wait on f_full, t_free, a_ready, total.
#ifdef C5_STATS
__device__ unsigned long long g_c5_stats[8];
#define C5_CLK() clock64()
#define C5_STAT_ADD(i, v) atomicAdd(&g_c5_stats[i], (unsigned long long)(v))
extern "C" int ngp_c5_stats(unsigned long long *out8, int reset) {
    if (out8 && cudaMemcpyFromSymbol(out8, g_c5_stats, sizeof(g_c5_stats)) != cudaSuccess) return 1;
    if (reset) {
        unsigned long long z[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        if (cudaMemcpyToSymbol(g_c5_stats, z, sizeof(z)) != cudaSuccess) return 1;
    }
    return 0;
}
#else
#define C5_CLK() 0ll
#define C5_STAT_ADD(i, v) ((void)(v))
#endif

struct __align__(128) C5FwdSmem {
    C5FwdWeights w;
    FqLevel lv[FQ_L];
    C5Bars bars[2];
    uint32_t tmem;
    int all_safe;
};

struct C5FwdParams {
    const float *x, *dirs, *table;
    float bound;
    int align_corners;
    float *sigma, *rgb;
    uint64_t *masks;
    int64_t B;
};

// ------------------------------------------------------------------------------------------------ forward
template <bool SAFE>
__device__ __forceinline__ void c5_forward_gather(C5FwdSmem &S, const C5FwdParams &P, uint32_t tmem, int warp, int lane) {
    const int s = warp >> 2, q = warp & 3, row = 32 * q + lane;
    const uint32_t lane_addr = tmem + s * C5_SLOT_COLS + ((uint32_t)(32 * q) << 16);
    const uint32_t bar_full = tc5::smem_u32(&S.bars[s].f_full), bar_empty = tc5::smem_u32(&S.bars[s].f_empty);
    const int64_t ntiles = (P.B + C5_ROWS - 1) / C5_ROWS;
    long long st_wait = 0;
    const long long st_t0 = C5_CLK();
    for (int64_t it = 0;; ++it) {
        const int64_t tile = blockIdx.x + (int64_t)gridDim.x * (2 * it + s);
        if (tile >= ntiles) break;
        const int64_t b = tile * C5_ROWS + row;
        float x01[3];
        const bool valid = fq_load_point(P.x, b, P.B, P.bound, x01);
#ifdef C5_PREFETCH
        // Experiment, measured and NOT the default: the next pair's 16 gathers in flight while the current pair is
        // interpolated.  At the 128 registers per thread that 448 threads get, it spills 116 bytes and the estimator's
        // forward goes 20.5 -> 22.6 ms (7.9x10^7 ray-ordered samples): more loads per gather thread is not what this
        // kernel is short of.
        FqRaw raw[2][2];
        float frac[2][2][3];
        c5_pair_issue<SAFE>(P.table, S.lv, 0, x01, P.align_corners, raw[0], frac[0]);
#pragma unroll
        for (int lp = 0; lp < 8; ++lp) {
            if (lp + 1 < 8) c5_pair_issue<SAFE>(P.table, S.lv, lp + 1, x01, P.align_corners, raw[(lp + 1) & 1], frac[(lp + 1) & 1]);
            float feat[4];
            c5_pair_consume(raw[lp & 1], frac[lp & 1], valid, feat);
            if (lp == 0) {
                if (it > 0) tc5::mbar_wait_wd(bar_empty, (uint32_t)((it - 1) & 1));
                tc5::fence_after_sync();
            }
            c5_put_features(lane_addr, lp, feat);
        }
#else
#pragma unroll 1
        for (int lp = 0; lp < 8; ++lp) {
            float feat[4];
            c5_encode_pair<SAFE>(P.table, S.lv, lp, x01, valid, P.align_corners, feat);
            if (lp == 0) {
                // the previous tile's first-layer products have read the feature columns (long done unless the chain is behind)
                const long long st_a = C5_CLK();
                if (it > 0) tc5::mbar_wait_wd(bar_empty, (uint32_t)((it - 1) & 1));
                st_wait += C5_CLK() - st_a;
                tc5::fence_after_sync();
            }
            c5_put_features(lane_addr, lp, feat);
        }
#endif
        tc5::tmem_st_wait();
        tc5::fence_before_sync();
        tc5::mbar_arrive(bar_full);
    }
    if (lane == 0) {
        C5_STAT_ADD(0, st_wait);
        C5_STAT_ADD(1, C5_CLK() - st_t0);
    }
}

// Chain warps: the five epilogues of a tile, alternating between the two slots step by step so that one slot's epilogue
// runs while the other slot's next layer is on the tensor core.  No state is carried between steps (per-row inputs are
// re-read from global memory before the barrier wait, outputs are written as soon as they exist), so each step's code
// exists once with the slot as a run-time index.  d_full completes 5 times per tile: wait k of tile `it` has parity
// (5 it + k) & 1.
template <bool MASK>
__device__ __forceinline__ void c5_forward_chain(C5FwdSmem &S, const C5FwdParams &P, uint32_t tmem, int warp, int lane) {
    const int q = warp & 3, row = 32 * q + lane;
    const int64_t ntiles = (P.B + C5_ROWS - 1) / C5_ROWS;
    const uint32_t lane_base = tmem + ((uint32_t)(32 * q) << 16);
    long long st_wait = 0;
    const long long st_t0 = C5_CLK();
    for (int64_t it = 0;; ++it) {
        const int64_t tile0 = blockIdx.x + (int64_t)gridDim.x * (2 * it);
        const int n_act = tile0 >= ntiles ? 0 : (tile0 + gridDim.x >= ntiles ? 1 : 2);
        if (n_act == 0) break;
        const uint32_t par = (uint32_t)(5 * it);
#define C5_SLOT_VARS                                                                            \
        const uint32_t lane_addr = lane_base + s * C5_SLOT_COLS;                                \
        const uint32_t bar_d = tc5::smem_u32(&S.bars[s].d_full), bar_a = tc5::smem_u32(&S.bars[s].a_ready); \
        const int64_t b = (tile0 + (int64_t)gridDim.x * s) * C5_ROWS + row;                     \
        const bool inb = b < P.B;
        // ---- sigma hidden
#pragma unroll 1
        for (int s = 0; s < n_act; ++s) {
            C5_SLOT_VARS
            { const long long st_a = C5_CLK(); tc5::mbar_wait_wd(bar_d, (par + 0u) & 1u); st_wait += C5_CLK() - st_a; }
            tc5::fence_after_sync();
            const uint64_t m = c5_relu_to_a<MASK>(lane_addr);
            c5_chain_arrive(bar_a);
            if (MASK && inb) P.masks[b * 3] = m;
        }
        // ---- sigma-net output o (column 0 = log sigma, 1..15 = geo_feat) -> colour input [SH(d) | 0, geo_feat]
#pragma unroll 1
        for (int s = 0; s < n_act; ++s) {
            C5_SLOT_VARS
            float dir[3];
#pragma unroll
            for (int a = 0; a < 3; ++a) dir[a] = inb ? __ldg(P.dirs + b * 3 + a) : 0.0f;
            { const long long st_a = C5_CLK(); tc5::mbar_wait_wd(bar_d, (par + 1u) & 1u); st_wait += C5_CLK() - st_a; }
            tc5::fence_after_sync();
            float o[16];
            c5_get16(lane_addr, 0, o);
            const float log_sigma = o[0];
            o[0] = 0.0f;
            c5_put16(lane_addr, 16, o);
            float sh[16];
            ngp_sh4(dir[0], dir[1], dir[2], sh);
            c5_put16(lane_addr, 0, sh);
            c5_chain_arrive(bar_a);
            if (inb) P.sigma[b] = expf(log_sigma);  // trunc_exp forward (activation.py:9-11)
        }
        // ---- colour hidden 1, 2
#pragma unroll 1
        for (int k = 0; k < 2; ++k) {
#pragma unroll 1
            for (int s = 0; s < n_act; ++s) {
                C5_SLOT_VARS
                { const long long st_a = C5_CLK(); tc5::mbar_wait_wd(bar_d, (par + 2u + (uint32_t)k) & 1u); st_wait += C5_CLK() - st_a; }
                tc5::fence_after_sync();
                const uint64_t m = c5_relu_to_a<MASK>(lane_addr);
                c5_chain_arrive(bar_a);
                if (MASK && inb) P.masks[b * 3 + 1 + k] = m;
            }
        }
        // ---- colour logits
#pragma unroll 1
        for (int s = 0; s < n_act; ++s) {
            C5_SLOT_VARS
            (void)bar_a;
            { const long long st_a = C5_CLK(); tc5::mbar_wait_wd(bar_d, (par + 4u) & 1u); st_wait += C5_CLK() - st_a; }
            tc5::fence_after_sync();
            uint32_t r[4];
            tc5::tmem_ld4_nowait(lane_addr + C5_D, r);
            tc5::tmem_ld_wait();
            tc5::fence_before_sync();
            tc5::mbar_arrive(tc5::smem_u32(&S.bars[s].t_free));
            if (inb) {
#pragma unroll
                for (int c = 0; c < 3; ++c) P.rgb[b * 3 + c] = 1.0f / (1.0f + expf(-__uint_as_float(r[c])));
            }
        }
    }
    if (lane == 0) {
        C5_STAT_ADD(2, st_wait);
        C5_STAT_ADD(3, C5_CLK() - st_t0);
    }
}

// MMA thread of slot s: the five layers of each tile.  a_ready completes 4 times per tile.
__device__ __forceinline__ void c5_forward_mma(C5FwdSmem &S, const C5FwdParams &P, uint32_t tmem, int s) {
    const int64_t ntiles = (P.B + C5_ROWS - 1) / C5_ROWS;
    const uint32_t t = tmem + s * C5_SLOT_COLS;
    const uint32_t bar_ff = tc5::smem_u32(&S.bars[s].f_full), bar_fe = tc5::smem_u32(&S.bars[s].f_empty),
                   bar_d = tc5::smem_u32(&S.bars[s].d_full), bar_a = tc5::smem_u32(&S.bars[s].a_ready),
                   bar_t = tc5::smem_u32(&S.bars[s].t_free);
    const uint32_t s0h = tc5::smem_u32(S.w.s0[0]), s0l = tc5::smem_u32(S.w.s0[1]), s1h = tc5::smem_u32(S.w.s1[0]),
                   s1l = tc5::smem_u32(S.w.s1[1]), c0h = tc5::smem_u32(S.w.c0[0]), c0l = tc5::smem_u32(S.w.c0[1]),
                   c1h = tc5::smem_u32(S.w.c1[0]), c1l = tc5::smem_u32(S.w.c1[1]), c2h = tc5::smem_u32(S.w.c2[0]),
                   c2l = tc5::smem_u32(S.w.c2[1]);
    long long st_ff = 0, st_t = 0, st_a_ = 0;
    const long long st_t0 = C5_CLK();
    for (int64_t it = 0;; ++it) {
        const int64_t tile = blockIdx.x + (int64_t)gridDim.x * (2 * it + s);
        if (tile >= ntiles) break;
        const uint32_t pa = (uint32_t)(4 * it);
        { const long long c = C5_CLK(); tc5::mbar_wait_wd(bar_ff, (uint32_t)(it & 1)); st_ff += C5_CLK() - c; }
        { const long long c = C5_CLK(); if (it > 0) tc5::mbar_wait_wd(bar_t, (uint32_t)((it - 1) & 1)); st_t += C5_CLK() - c; }
        tc5::fence_after_sync();
        c5_issue_layer<64, 32>(t + C5_D, t + C5_FHI, t + C5_FLO, s0h, s0l);   // sigma L1: features -> 64
        tc5::mma_commit(bar_fe);
        tc5::mma_commit(bar_d);
        { const long long c = C5_CLK(); tc5::mbar_wait_wd(bar_a, (pa + 0u) & 1u); st_a_ += C5_CLK() - c; }
        tc5::fence_after_sync();
        c5_issue_layer<16, 64>(t + C5_D, t + C5_HHI, t + C5_HLO, s1h, s1l);   // sigma L2: 64 -> 16
        tc5::mma_commit(bar_d);
        { const long long c = C5_CLK(); tc5::mbar_wait_wd(bar_a, (pa + 1u) & 1u); st_a_ += C5_CLK() - c; }
        tc5::fence_after_sync();
        c5_issue_layer<64, 32>(t + C5_D, t + C5_HHI, t + C5_HLO, c0h, c0l);   // colour L1: [SH | o] (32) -> 64
        tc5::mma_commit(bar_d);
        { const long long c = C5_CLK(); tc5::mbar_wait_wd(bar_a, (pa + 2u) & 1u); st_a_ += C5_CLK() - c; }
        tc5::fence_after_sync();
        c5_issue_layer<64, 64>(t + C5_D, t + C5_HHI, t + C5_HLO, c1h, c1l);   // colour L2: 64 -> 64
        tc5::mma_commit(bar_d);
        { const long long c = C5_CLK(); tc5::mbar_wait_wd(bar_a, (pa + 3u) & 1u); st_a_ += C5_CLK() - c; }
        tc5::fence_after_sync();
        c5_issue_layer<16, 64>(t + C5_D, t + C5_HHI, t + C5_HLO, c2h, c2l);   // colour L3: 64 -> 3 (16)
        tc5::mma_commit(bar_d);
    }
    C5_STAT_ADD(4, st_ff);
    C5_STAT_ADD(5, st_t);
    C5_STAT_ADD(6, st_a_);
    C5_STAT_ADD(7, C5_CLK() - st_t0);
}

__device__ __forceinline__ void c5_init_bars(C5Bars *bars) {
    for (int s = 0; s < 2; ++s) {
        tc5::mbar_init(tc5::smem_u32(&bars[s].f_full), 128);
        tc5::mbar_init(tc5::smem_u32(&bars[s].f_empty), 1);
        tc5::mbar_init(tc5::smem_u32(&bars[s].d_full), 1);
        tc5::mbar_init(tc5::smem_u32(&bars[s].a_ready), 128);
        tc5::mbar_init(tc5::smem_u32(&bars[s].t_free), 128);
        tc5::mbar_init(tc5::smem_u32(&bars[s].g_full), 1);
    }
    tc5::fence_mbar_init();
}

__global__ void __launch_bounds__(C5_THREADS, 1)
k_cfield5_forward(C5FwdParams P, const int32_t *__restrict__ offsets, float S_, uint32_t H, uint32_t gridtype,
                  const float *__restrict__ Ws0, const float *__restrict__ Ws1, const float *__restrict__ Wc0,
                  const float *__restrict__ Wc1, const float *__restrict__ Wc2) {
    extern __shared__ __align__(128) unsigned char c5_smem_raw[];
    C5FwdSmem &S = *reinterpret_cast<C5FwdSmem *>(c5_smem_raw);
    // warp index and (below) the TMEM base as values the compiler can PROVE warp-uniform (a broadcast shuffle, the
    // canonical_warp_idx_sync idiom): tcgen05.mma takes its TMEM addresses and descriptors from uniform registers, and
    // for operands it cannot prove uniform the compiler wraps every MMA in an ELECT / R2UR.BROADCAST / branch loop
    // (~100 cycles per MMA for the issuing thread, 96 MMAs per tile: scripts/c5_roles.py had the MMA threads issuing
    // 58 % of the time and never waiting for features)
    const int tid = threadIdx.x, warp = __shfl_sync(0xffffffffu, tid >> 5, 0), lane = tid & 31;
    fq_stage_levels(S.lv, &S.all_safe, offsets, S_, H, gridtype, P.align_corners);
    c5_stage_forward_weights(S.w, Ws0, Ws1, Wc0, Wc1, Wc2);
    if (tid == 0) c5_init_bars(S.bars);
    if (warp == C5_MMA_WARP) tc5::tmem_alloc(tc5::smem_u32(&S.tmem), 512);
    tc5::fence_async_smem();
    tc5::fence_before_sync();
    __syncthreads();
    tc5::fence_after_sync();
    const uint32_t tmem = __shfl_sync(0xffffffffu, S.tmem, 0);

    if (warp < C5_GATHER_WARPS) {
        if (S.all_safe) c5_forward_gather<true>(S, P, tmem, warp, lane);
        else c5_forward_gather<false>(S, P, tmem, warp, lane);
    } else if (warp < C5_MMA_WARP) {
        if (P.masks) c5_forward_chain<true>(S, P, tmem, warp, lane);
        else c5_forward_chain<false>(S, P, tmem, warp, lane);
    } else {
        if (lane == 0) c5_forward_mma(S, P, tmem, warp - C5_MMA_WARP);
        __syncwarp();
    }
    tc5::fence_before_sync();
    __syncthreads();
    if (warp == C5_MMA_WARP) tc5::tmem_dealloc(tmem, 512);
}

// ------------------------------------------------------------------------------------------------ backward to (x, d)
// Every product is (row gradient) x W: the B tile of a layer is W^T, B(n, k) = W[k-th downstream unit][n-th upstream unit].
struct __align__(128) C5BwdWeights {
    float c2[2][64 * 8];   // dHc2 = dRGB  W_c2   [64 x 8]  (K = 3 -> 8)
    float c1[2][64 * 64];  // dHc1 = dHc2m W_c1   [64 x 64]
    float c0[2][32 * 64];  // dIn  = dHc1m W_c0   [32 x 64]: n < 16 SH, n >= 16 sigma-net output n-16 (n = 16: zero)
    float s1[2][64 * 16];  // dHs  = dO    W_s1   [64 x 16]
    float s0[2][32 * 64];  // dFeat = dHsm W_s0   [32 x 64]
};
struct __align__(128) C5BwdSmem {
    C5BwdWeights w;
    FqLevel lv[FQ_L];
    C5Bars bars[2];
    uint32_t tmem;
    int all_safe;
};
struct C5BwdParams {
    const float *x, *dirs, *table;
    float bound;
    int align_corners;
    const float *sigma, *rgb;
    const uint64_t *masks;
    const float *grad_sigma, *grad_rgb;
    float *grad_x, *grad_d;
    int64_t B;
};

// Chain warps, backward: five A-operand steps per tile, alternating between the slots like the forward.  d_full completes
// 4 times per tile (parity of wait k = k & 1), the fifth product signals the gather warps (g_full).
__device__ __forceinline__ void c5_backward_chain(C5BwdSmem &S, const C5BwdParams &P, uint32_t tmem, int warp, int lane) {
    const int q = warp & 3, row = 32 * q + lane;
    const int64_t ntiles = (P.B + C5_ROWS - 1) / C5_ROWS;
    const float e_lo = expf(-15.0f), e_hi = expf(15.0f);
    const uint32_t lane_base = tmem + ((uint32_t)(32 * q) << 16);
    const unsigned long long *masks = reinterpret_cast<const unsigned long long *>(P.masks);
    for (int64_t it = 0;; ++it) {
        const int64_t tile0 = blockIdx.x + (int64_t)gridDim.x * (2 * it);
        const int n_act = tile0 >= ntiles ? 0 : (tile0 + gridDim.x >= ntiles ? 1 : 2);
        if (n_act == 0) break;
        // ---- dRGB = dL/drgb .* sigmoid'
#pragma unroll 1
        for (int s = 0; s < n_act; ++s) {
            C5_SLOT_VARS
            (void)bar_d;
            float a0[16];
#pragma unroll
            for (int i = 0; i < 16; ++i) a0[i] = 0.0f;
            if (inb) {
#pragma unroll
                for (int c = 0; c < 3; ++c) {
                    const float cv = __ldg(P.rgb + b * 3 + c);
                    a0[c] = __ldg(P.grad_rgb + b * 3 + c) * cv * (1.0f - cv);
                }
            }
            // the slot is free once the gather warps have taken the previous tile's feature gradients
            if (it > 0) tc5::mbar_wait_wd(tc5::smem_u32(&S.bars[s].t_free), (uint32_t)((it - 1) & 1));
            tc5::fence_after_sync();
            c5_put16(lane_addr, 0, a0);  // K = 8 of these 16 columns are read
            c5_chain_arrive(bar_a);
        }
        // ---- dHc2 .* relu', dHc1 .* relu'
#pragma unroll 1
        for (int k = 0; k < 2; ++k) {
#pragma unroll 1
            for (int s = 0; s < n_act; ++s) {
                C5_SLOT_VARS
                const uint64_t m = inb ? __ldg(masks + b * 3 + 2 - k) : 0ull;
                tc5::mbar_wait_wd(bar_d, (uint32_t)k & 1u);
                tc5::fence_after_sync();
                c5_mask_to_a(lane_addr, m);
                c5_chain_arrive(bar_a);
            }
        }
        // ---- dIn = [dSH 16 | dO 16]: dO (+ the sigma gradient on column 0) goes on; SH backward to the direction
#pragma unroll 1
        for (int s = 0; s < n_act; ++s) {
            C5_SLOT_VARS
            float dir[3] = {0.0f, 0.0f, 0.0f}, go0 = 0.0f;
            if (inb) {
                go0 = __ldg(P.grad_sigma + b) * fminf(fmaxf(__ldg(P.sigma + b), e_lo), e_hi);  // trunc_exp backward (activation.py:13-17)
#pragma unroll
                for (int c = 0; c < 3; ++c) dir[c] = __ldg(P.dirs + b * 3 + c);
            }
            tc5::mbar_wait_wd(bar_d, 0u);
            tc5::fence_after_sync();
            float dout[16], dsh[16];
            c5_get16(lane_addr, 16, dout);
            c5_get16(lane_addr, 0, dsh);
            dout[0] += go0;  // its colour weight is zero, so the product left 0 there
            c5_put16(lane_addr, 0, dout);
            c5_chain_arrive(bar_a);
            if (P.grad_d) {
                float jx[16], jy[16], jz[16];
                ngp_sh4_jac(dir[0], dir[1], dir[2], jx, jy, jz);
                float g0 = 0.0f, g1 = 0.0f, g2 = 0.0f;
#pragma unroll
                for (int i = 0; i < 16; ++i) {
                    g0 = fmaf(dsh[i], jx[i], g0);
                    g1 = fmaf(dsh[i], jy[i], g1);
                    g2 = fmaf(dsh[i], jz[i], g2);
                }
                if (inb) {
                    P.grad_d[b * 3] = g0;
                    P.grad_d[b * 3 + 1] = g1;
                    P.grad_d[b * 3 + 2] = g2;
                }
            }
        }
        // ---- dHs .* relu'
#pragma unroll 1
        for (int s = 0; s < n_act; ++s) {
            C5_SLOT_VARS
            const uint64_t m = inb ? __ldg(masks + b * 3) : 0ull;
            tc5::mbar_wait_wd(bar_d, 1u);
            tc5::fence_after_sync();
            c5_mask_to_a(lane_addr, m);
            c5_chain_arrive(bar_a);
        }
    }
}

// MMA thread of slot s, backward.  a_ready completes 5 times per tile.
__device__ __forceinline__ void c5_backward_mma(C5BwdSmem &S, const C5BwdParams &P, uint32_t tmem, int s) {
    const int64_t ntiles = (P.B + C5_ROWS - 1) / C5_ROWS;
    const uint32_t t = tmem + s * C5_SLOT_COLS;
    const uint32_t bar_d = tc5::smem_u32(&S.bars[s].d_full), bar_a = tc5::smem_u32(&S.bars[s].a_ready),
                   bar_g = tc5::smem_u32(&S.bars[s].g_full);
    const uint32_t c2h = tc5::smem_u32(S.w.c2[0]), c2l = tc5::smem_u32(S.w.c2[1]), c1h = tc5::smem_u32(S.w.c1[0]),
                   c1l = tc5::smem_u32(S.w.c1[1]), c0h = tc5::smem_u32(S.w.c0[0]), c0l = tc5::smem_u32(S.w.c0[1]),
                   s1h = tc5::smem_u32(S.w.s1[0]), s1l = tc5::smem_u32(S.w.s1[1]), s0h = tc5::smem_u32(S.w.s0[0]),
                   s0l = tc5::smem_u32(S.w.s0[1]);
    for (int64_t it = 0;; ++it) {
        const int64_t tile = blockIdx.x + (int64_t)gridDim.x * (2 * it + s);
        if (tile >= ntiles) break;
        const uint32_t pa = (uint32_t)(5 * it);
        tc5::mbar_wait_wd(bar_a, (pa + 0u) & 1u);
        tc5::fence_after_sync();
        c5_issue_layer<64, 8>(t + C5_D, t + C5_HHI, t + C5_HLO, c2h, c2l);    // dHc2 = dRGB W_c2
        tc5::mma_commit(bar_d);
        tc5::mbar_wait_wd(bar_a, (pa + 1u) & 1u);
        tc5::fence_after_sync();
        c5_issue_layer<64, 64>(t + C5_D, t + C5_HHI, t + C5_HLO, c1h, c1l);   // dHc1 = dHc2m W_c1
        tc5::mma_commit(bar_d);
        tc5::mbar_wait_wd(bar_a, (pa + 2u) & 1u);
        tc5::fence_after_sync();
        c5_issue_layer<32, 64>(t + C5_D, t + C5_HHI, t + C5_HLO, c0h, c0l);   // dIn = dHc1m W_c0
        tc5::mma_commit(bar_d);
        tc5::mbar_wait_wd(bar_a, (pa + 3u) & 1u);
        tc5::fence_after_sync();
        c5_issue_layer<64, 16>(t + C5_D, t + C5_HHI, t + C5_HLO, s1h, s1l);   // dHs = dO W_s1
        tc5::mma_commit(bar_d);
        tc5::mbar_wait_wd(bar_a, (pa + 4u) & 1u);
        tc5::fence_after_sync();
        c5_issue_layer<32, 64>(t + C5_D, t + C5_HHI, t + C5_HLO, s0h, s0l);   // dFeat = dHsm W_s0
        tc5::mma_commit(bar_g);
    }
}

template <bool SAFE>
__device__ __forceinline__ void c5_backward_gather(C5BwdSmem &S, const C5BwdParams &P, uint32_t tmem, int warp, int lane) {
    const int s = warp >> 2, q = warp & 3, row = 32 * q + lane;
    const uint32_t lane_addr = tmem + s * C5_SLOT_COLS + ((uint32_t)(32 * q) << 16);
    const uint32_t bar_g = tc5::smem_u32(&S.bars[s].g_full), bar_free = tc5::smem_u32(&S.bars[s].t_free);
    const int64_t ntiles = (P.B + C5_ROWS - 1) / C5_ROWS;
    const float den = 2.0f * P.bound;
    for (int64_t it = 0;; ++it) {
        const int64_t tile = blockIdx.x + (int64_t)gridDim.x * (2 * it + s);
        if (tile >= ntiles) break;
        const int64_t b = tile * C5_ROWS + row;
        float x01[3], gx[3] = {0.0f, 0.0f, 0.0f};
        const bool valid = fq_load_point(P.x, b, P.B, P.bound, x01);
        tc5::mbar_wait_wd(bar_g, (uint32_t)(it & 1));
        tc5::fence_after_sync();
        // move the thread's 32 feature gradients to its private columns and hand the accumulator back to the chain
#pragma unroll
        for (int c = 0; c < 2; ++c) {
            uint32_t r[16];
            tc5::tmem_ld16_nowait(lane_addr + C5_D + 16 * c, r);
            tc5::tmem_ld_wait();
            if (!valid) {
#pragma unroll
                for (int i = 0; i < 16; ++i) r[i] = 0u;
            }
            tc5::tmem_st16(lane_addr + C5_DF + 16 * c, r);
        }
        tc5::tmem_st_wait();
        tc5::fence_before_sync();
        tc5::mbar_arrive(bar_free);
#pragma unroll 1
        for (int lp = 0; lp < 8; ++lp) {
            uint32_t r[4];
            tc5::tmem_ld4_nowait(lane_addr + C5_DF + 4 * lp, r);
            tc5::tmem_ld_wait();
            const float df[4] = {__uint_as_float(r[0]), __uint_as_float(r[1]), __uint_as_float(r[2]), __uint_as_float(r[3])};
            c5_encode_pair_backward<SAFE>(P.table, S.lv, lp, x01, P.align_corners, df, gx);
        }
        if (b < P.B) {
#pragma unroll
            for (int a = 0; a < 3; ++a) P.grad_x[b * 3 + a] = __fdiv_rn(gx[a], den);
        }
    }
}

__global__ void __launch_bounds__(C5_THREADS, 1)
k_cfield5_backward(C5BwdParams P, const int32_t *__restrict__ offsets, float S_, uint32_t H, uint32_t gridtype,
                   const float *__restrict__ Ws0, const float *__restrict__ Ws1, const float *__restrict__ Wc0,
                   const float *__restrict__ Wc1, const float *__restrict__ Wc2) {
    extern __shared__ __align__(128) unsigned char c5_smem_raw[];
    C5BwdSmem &S = *reinterpret_cast<C5BwdSmem *>(c5_smem_raw);
    const int tid = threadIdx.x, warp = __shfl_sync(0xffffffffu, tid >> 5, 0), lane = tid & 31;
    fq_stage_levels(S.lv, &S.all_safe, offsets, S_, H, gridtype, P.align_corners);
    c5_stage_tile(S.w.c2[0], S.w.c2[1], 64, 8, [&](int n, int k) { return k < 3 ? Wc2[k * FQ_HID + n] : 0.0f; });
    c5_stage_tile(S.w.c1[0], S.w.c1[1], 64, 64, [&](int n, int k) { return Wc1[k * FQ_HID + n]; });
    c5_stage_tile(S.w.c0[0], S.w.c0[1], 32, 64, [&](int n, int k) {
        return n < 16 ? Wc0[k * 31 + n] : (n == 16 ? 0.0f : Wc0[k * 31 + n - 1]);
    });
    c5_stage_tile(S.w.s1[0], S.w.s1[1], 64, 16, [&](int n, int k) { return Ws1[k * FQ_HID + n]; });
    c5_stage_tile(S.w.s0[0], S.w.s0[1], 32, 64, [&](int n, int k) { return Ws0[k * FQ_F + n]; });
    if (tid == 0) c5_init_bars(S.bars);
    if (warp == C5_MMA_WARP) tc5::tmem_alloc(tc5::smem_u32(&S.tmem), 512);
    tc5::fence_async_smem();
    tc5::fence_before_sync();
    __syncthreads();
    tc5::fence_after_sync();
    const uint32_t tmem = __shfl_sync(0xffffffffu, S.tmem, 0);

    if (warp < C5_GATHER_WARPS) {
        if (S.all_safe) c5_backward_gather<true>(S, P, tmem, warp, lane);
        else c5_backward_gather<false>(S, P, tmem, warp, lane);
    } else if (warp < C5_MMA_WARP) {
        c5_backward_chain(S, P, tmem, warp, lane);
    } else {
        if (lane == 0) c5_backward_mma(S, P, tmem, warp - C5_MMA_WARP);
        __syncwarp();
    }
    tc5::fence_before_sync();
    __syncthreads();
    if (warp == C5_MMA_WARP) tc5::tmem_dealloc(tmem, 512);
}

// ------------------------------------------------------------------------------------------------ launchers
// one CTA per SM: a CTA owns all 512 TMEM columns, so a second resident CTA would only block in tcgen05.alloc
static int c5_smem_bytes(size_t need) { return (int)(need > 120 * 1024 ? need : 120 * 1024); }

static int c5_grid(int64_t B) {
    int dev = 0, sms = NGP_SM_COUNT;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const int64_t tiles = ngp_div_up(B, C5_ROWS);
    return (int)(tiles < sms ? tiles : sms);
}

int c5_field_color_forward(const float *xyzs, const float *dirs, float bound, const float *embeddings,
                           const int32_t *offsets, float S, uint32_t H, uint32_t gridtype, int align_corners,
                           const float *Ws0, const float *Ws1, const float *Wc0, const float *Wc1, const float *Wc2,
                           float *sigma, float *rgb, void *masks, int64_t M, cudaStream_t stream) {
    const int smem = c5_smem_bytes(sizeof(C5FwdSmem));
    if (cudaFuncSetAttribute(k_cfield5_forward, cudaFuncAttributeMaxDynamicSharedMemorySize, smem) != cudaSuccess) {
        ngp_set_error("ngp_field_color_forward: cannot reserve %d B of shared memory", smem);
        return NGP_ERR_CUDA;
    }
    C5FwdParams P;
    P.x = xyzs; P.dirs = dirs; P.table = embeddings; P.bound = bound; P.align_corners = align_corners;
    P.sigma = sigma; P.rgb = rgb; P.masks = (uint64_t *)masks; P.B = M;
    k_cfield5_forward<<<c5_grid(M), C5_THREADS, smem, stream>>>(P, offsets, S, H, gridtype, Ws0, Ws1, Wc0, Wc1, Wc2);
    NGP_CHECK_LAUNCH("ngp_field_color_forward (tcgen05)");
    return NGP_OK;
}

int c5_field_color_backward(const float *xyzs, const float *dirs, float bound, const float *embeddings,
                            const int32_t *offsets, float S, uint32_t H, uint32_t gridtype, int align_corners,
                            const float *Ws0, const float *Ws1, const float *Wc0, const float *Wc1, const float *Wc2,
                            const float *sigma, const float *rgb, const void *masks, const float *grad_sigma,
                            const float *grad_rgb, float *grad_xyzs, float *grad_dirs, int64_t M, cudaStream_t stream) {
    const int smem = c5_smem_bytes(sizeof(C5BwdSmem));
    if (cudaFuncSetAttribute(k_cfield5_backward, cudaFuncAttributeMaxDynamicSharedMemorySize, smem) != cudaSuccess) {
        ngp_set_error("ngp_field_color_backward: cannot reserve %d B of shared memory", smem);
        return NGP_ERR_CUDA;
    }
    C5BwdParams P;
    P.x = xyzs; P.dirs = dirs; P.table = embeddings; P.bound = bound; P.align_corners = align_corners;
    P.sigma = sigma; P.rgb = rgb; P.masks = (const uint64_t *)masks; P.grad_sigma = grad_sigma; P.grad_rgb = grad_rgb;
    P.grad_x = grad_xyzs; P.grad_d = grad_dirs; P.B = M;
    k_cfield5_backward<<<c5_grid(M), C5_THREADS, smem, stream>>>(P, offsets, S, H, gridtype, Ws0, Ws1, Wc0, Wc1, Wc2);
    NGP_CHECK_LAUNCH("ngp_field_color_backward (tcgen05)");
    return NGP_OK;
}
