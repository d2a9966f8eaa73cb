// dfield5.cu — warp-specialised tcgen05 kernels of the fused DENSITY field query, the headline path of the planner:
// NeRFNetwork.density (nerf/network.py:126-143) = hash-grid encode (gridencoder.cu:75-178) -> sigma net 32-64-16
// (network.py:33-47) -> trunc_exp (activation.py:5-17), and its backward to the points (the collision cost's dL/dx).
// Same machine mapping as cfield5.cu (building blocks in cfield5.cuh): 128-point tiles = the 128 TMEM lanes, gather
// threads write the interpolated features straight into TMEM as TF32 (hi, lo) pairs, one thread per tile slot issues the
// two layers as tcgen05.mma kind::tf32 (3xTF32, weights pre-split in shared memory, accumulators in TMEM), a chain
// warpgroup runs the epilogues with tcgen05.ld / tcgen05.st; two tile slots per SM so that the gathers of one tile
// overlap the layers of the other.  These are the default implementation behind ngp_field_query_forward / backward
// (field.cu keeps the mma.sync kernels as the cross-check: ngp_set_field_impl(0) / NGP_B200_FIELD_IMPL=mma).
#include "cfield5.cuh"

// -DD5_STATS=1 (dev aid, scripts/build_variant.sh): block 0 prints where the roles of the value+grad kernel spend their cycles
#ifndef D5_STATS
#define D5_STATS 0
#endif
#if D5_STATS
#define D5_TIMED(acc, stmt) { const long long t_ = clock64(); stmt; acc += clock64() - t_; }
#else
#define D5_TIMED(acc, stmt) { stmt; }
#endif

struct __align__(128) D5Weights {
    float s0[2][64 * 32];  // forward: sigma L1 [64 x 32]            backward: dFeat = dHm W_s0  [32 x 64]
    float s1[2][16 * 64];  // forward: sigma L2 [16 x 64]            backward: dH    = dO  W_s1  [64 x 16]
};
struct __align__(128) D5Smem {
    D5Weights w;
    FqLevel lv[FQ_L];
    C5Bars bars[2];
    uint32_t tmem;
    int all_safe;
    __align__(128) float stage[4][2][32 * 15];  // forward: geo_feat rows of one chain warp (1920 B = 32 rows), double-buffered for the bulk store
};

struct D5FwdParams {
    const float *x, *table;
    float bound;
    int align_corners;
    float *sigma, *geo;
    unsigned long long *masks;
    int64_t B;
};
struct D5BwdParams {
    const float *x, *table;
    float bound;
    int align_corners;
    const float *sigma;
    const unsigned long long *masks;
    const float *grad_sigma, *grad_geo;
    float *grad_x;
    int64_t B;
};

__device__ __forceinline__ void d5_init_bars(C5Bars *bars, int gather_threads = 128) {
    for (int s = 0; s < 2; ++s) {
        tc5::mbar_init(tc5::smem_u32(&bars[s].f_full), gather_threads);
        tc5::mbar_init(tc5::smem_u32(&bars[s].f_empty), 1);
        tc5::mbar_init(tc5::smem_u32(&bars[s].d_full), 1);
        tc5::mbar_init(tc5::smem_u32(&bars[s].a_ready), 128);
        tc5::mbar_init(tc5::smem_u32(&bars[s].t_free), 128);
        tc5::mbar_init(tc5::smem_u32(&bars[s].g_full), 1);
    }
    tc5::fence_mbar_init();
}

// ------------------------------------------------------------------------------------------------ forward
template <bool SAFE>
__device__ __forceinline__ void d5_forward_gather(D5Smem &S, const D5FwdParams &P, uint32_t tmem, int warp, int lane) {
    const int s = warp >> 2, q = warp & 3, row = 32 * q + lane;
    const uint32_t lane_addr = tmem + s * C5_SLOT_COLS + ((uint32_t)(32 * q) << 16);
    const uint32_t bar_full = tc5::smem_u32(&S.bars[s].f_full), bar_empty = tc5::smem_u32(&S.bars[s].f_empty);
    const int64_t ntiles = (P.B + C5_ROWS - 1) / C5_ROWS;
    for (int64_t it = 0;; ++it) {
        const int64_t tile = blockIdx.x + (int64_t)gridDim.x * (2 * it + s);
        if (tile >= ntiles) break;
        const int64_t b = tile * C5_ROWS + row;
        float x01[3];
        const bool valid = fq_load_point(P.x, b, P.B, P.bound, x01);
#pragma unroll 1
        for (int lp = 0; lp < 8; ++lp) {
            float feat[4];
            c5_encode_pair<SAFE>(P.table, S.lv, lp, x01, valid, P.align_corners, feat);
            if (lp == 0) {
                if (it > 0) tc5::mbar_wait_wd(bar_empty, (uint32_t)((it - 1) & 1));
                tc5::fence_after_sync();
            }
            c5_put_features(lane_addr, lp, feat);
        }
        tc5::tmem_st_wait();
        tc5::fence_before_sync();
        tc5::mbar_arrive(bar_full);
    }
}

// chain warps: two epilogues per tile, alternating between the slots.  d_full completes twice per tile.
// geo_feat leaves through shared memory: a warp's 32 rows x 15 floats are 1920 contiguous bytes of the output, written by
// ONE bulk copy (cp.async.bulk shared -> global, the TMA unit) instead of 15 row-strided store instructions whose 32 lanes
// each cost an L1 wavefront — the unit that binds this kernel (the gathers need it).
__device__ __forceinline__ void d5_bulk_store(float *dst, const float *src_smem, uint32_t bytes) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst), "r"(tc5::smem_u32(src_smem)), "r"(bytes)
                 : "memory");
    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
}

template <bool MASK>
__device__ __forceinline__ void d5_forward_chain(D5Smem &S, const D5FwdParams &P, uint32_t tmem, int warp, int lane) {
    const int q = warp & 3, row = 32 * q + lane;
    const bool bulk_ok = P.geo && (reinterpret_cast<uintptr_t>(P.geo) & 15) == 0;
    uint32_t nbulk = 0;
    const int64_t ntiles = (P.B + C5_ROWS - 1) / C5_ROWS;
    const uint32_t lane_base = tmem + ((uint32_t)(32 * q) << 16);
    for (int64_t it = 0;; ++it) {
        const int64_t tile0 = blockIdx.x + (int64_t)gridDim.x * (2 * it);
        const int n_act = tile0 >= ntiles ? 0 : (tile0 + gridDim.x >= ntiles ? 1 : 2);
        if (n_act == 0) break;
        const uint32_t par = (uint32_t)(2 * it);
        // ---- hidden layer: ReLU (+ sign bits for the backward) -> A operand of the output layer
#pragma unroll 1
        for (int s = 0; s < n_act; ++s) {
            const uint32_t lane_addr = lane_base + s * C5_SLOT_COLS;
            const int64_t b = (tile0 + (int64_t)gridDim.x * s) * C5_ROWS + row;
            tc5::mbar_wait_wd(tc5::smem_u32(&S.bars[s].d_full), (par + 0u) & 1u);
            tc5::fence_after_sync();
            const uint64_t m = c5_relu_to_a<MASK>(lane_addr);
            c5_chain_arrive(tc5::smem_u32(&S.bars[s].a_ready));
            if (MASK && b < P.B) P.masks[b] = m;
        }
        // ---- output layer: column 0 = log sigma, 1..15 = geo_feat
#pragma unroll 1
        for (int s = 0; s < n_act; ++s) {
            const uint32_t lane_addr = lane_base + s * C5_SLOT_COLS;
            const int64_t b = (tile0 + (int64_t)gridDim.x * s) * C5_ROWS + row;
            tc5::mbar_wait_wd(tc5::smem_u32(&S.bars[s].d_full), (par + 1u) & 1u);
            tc5::fence_after_sync();
            float o[16];
            c5_get16(lane_addr, 0, o);
            tc5::fence_before_sync();
            tc5::mbar_arrive(tc5::smem_u32(&S.bars[s].t_free));
            if (b < P.B) P.sigma[b] = expf(o[0]);  // trunc_exp forward (activation.py:9-11)
            const int64_t b0 = b - lane;  // first row of this warp
            if (bulk_ok && b0 + 32 <= P.B) {
                float *st = S.stage[q][nbulk & 1];
                if (nbulk >= 2) {  // the copy that read this buffer two stores ago has drained it
                    if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
                    __syncwarp();
                }
#pragma unroll
                for (int i = 0; i < 15; ++i) st[lane * 15 + i] = o[1 + i];
                tc5::fence_async_smem();
                __syncwarp();
                if (lane == 0) d5_bulk_store(P.geo + b0 * 15, st, 32 * 15 * 4);
                ++nbulk;
            } else if (P.geo && b < P.B) {
#pragma unroll
                for (int i = 0; i < 15; ++i) P.geo[b * 15 + i] = o[1 + i];
            }
        }
    }
    if (lane == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");  // all bytes written before the kernel ends
}

__device__ __forceinline__ void d5_forward_mma(D5Smem &S, const D5FwdParams &P, uint32_t tmem, int s) {
    const int64_t ntiles = (P.B + C5_ROWS - 1) / C5_ROWS;
    const uint32_t t = tmem + s * C5_SLOT_COLS;
    const uint32_t bar_ff = tc5::smem_u32(&S.bars[s].f_full), bar_fe = tc5::smem_u32(&S.bars[s].f_empty),
                   bar_d = tc5::smem_u32(&S.bars[s].d_full), bar_a = tc5::smem_u32(&S.bars[s].a_ready),
                   bar_t = tc5::smem_u32(&S.bars[s].t_free);
    const uint32_t s0h = tc5::smem_u32(S.w.s0[0]), s0l = tc5::smem_u32(S.w.s0[1]), s1h = tc5::smem_u32(S.w.s1[0]),
                   s1l = tc5::smem_u32(S.w.s1[1]);
    for (int64_t it = 0;; ++it) {
        const int64_t tile = blockIdx.x + (int64_t)gridDim.x * (2 * it + s);
        if (tile >= ntiles) break;
        tc5::mbar_wait_wd(bar_ff, (uint32_t)(it & 1));
        if (it > 0) tc5::mbar_wait_wd(bar_t, (uint32_t)((it - 1) & 1));
        tc5::fence_after_sync();
        c5_issue_layer<64, 32>(t + C5_D, t + C5_FHI, t + C5_FLO, s0h, s0l);   // features -> 64
        tc5::mma_commit(bar_fe);
        tc5::mma_commit(bar_d);
        tc5::mbar_wait_wd(bar_a, (uint32_t)(it & 1));
        tc5::fence_after_sync();
        c5_issue_layer<16, 64>(t + C5_D, t + C5_HHI, t + C5_HLO, s1h, s1l);   // 64 -> 16
        tc5::mma_commit(bar_d);
    }
}

__global__ void __launch_bounds__(C5_THREADS, 1)
k_dfield5_forward(D5FwdParams P, const int32_t *__restrict__ offsets, float S_, uint32_t H, uint32_t gridtype,
                  const float *__restrict__ W0, const float *__restrict__ W1) {
    extern __shared__ __align__(128) unsigned char d5_smem_raw[];
    D5Smem &S = *reinterpret_cast<D5Smem *>(d5_smem_raw);
    const int tid = threadIdx.x, warp = __shfl_sync(0xffffffffu, tid >> 5, 0), lane = tid & 31;
    fq_stage_levels(S.lv, &S.all_safe, offsets, S_, H, gridtype, P.align_corners);
    c5_stage_tile(S.w.s0[0], S.w.s0[1], 64, 32, [&](int n, int k) { return W0[n * FQ_F + k]; });
    c5_stage_tile(S.w.s1[0], S.w.s1[1], 16, 64, [&](int n, int k) { return W1[n * FQ_HID + k]; });
    if (tid == 0) d5_init_bars(S.bars);
    if (warp == C5_MMA_WARP) tc5::tmem_alloc(tc5::smem_u32(&S.tmem), 512);
    tc5::fence_async_smem();
    tc5::fence_before_sync();
    __syncthreads();
    tc5::fence_after_sync();
    const uint32_t tmem = __shfl_sync(0xffffffffu, S.tmem, 0);
    if (warp < C5_GATHER_WARPS) {
        if (S.all_safe) d5_forward_gather<true>(S, P, tmem, warp, lane);
        else d5_forward_gather<false>(S, P, tmem, warp, lane);
    } else if (warp < C5_MMA_WARP) {
        if (P.masks) d5_forward_chain<true>(S, P, tmem, warp, lane);
        else d5_forward_chain<false>(S, P, tmem, warp, lane);
    } else {
        if (lane == 0) d5_forward_mma(S, P, tmem, warp - C5_MMA_WARP);
        __syncwarp();
    }
    tc5::fence_before_sync();
    __syncthreads();
    if (warp == C5_MMA_WARP) tc5::tmem_dealloc(tmem, 512);
}

// ------------------------------------------------------------------------------------------------ backward to x
// Every product is (row gradient) x W: the B tile of a layer is W^T (cfield5.cu).  a_ready completes twice per tile,
// d_full once, the feature-gradient product signals the gather warps (g_full).
__device__ __forceinline__ void d5_backward_chain(D5Smem &S, const D5BwdParams &P, uint32_t tmem, int warp, int lane) {
    const int q = warp & 3, row = 32 * q + lane;
    const int64_t ntiles = (P.B + C5_ROWS - 1) / C5_ROWS;
    const float e_lo = expf(-15.0f), e_hi = expf(15.0f);
    const uint32_t lane_base = tmem + ((uint32_t)(32 * q) << 16);
    for (int64_t it = 0;; ++it) {
        const int64_t tile0 = blockIdx.x + (int64_t)gridDim.x * (2 * it);
        const int n_act = tile0 >= ntiles ? 0 : (tile0 + gridDim.x >= ntiles ? 1 : 2);
        if (n_act == 0) break;
        // ---- dO = [dL/dsigma * exp(clamp(log sigma)), dL/dgeo]
#pragma unroll 1
        for (int s = 0; s < n_act; ++s) {
            const uint32_t lane_addr = lane_base + s * C5_SLOT_COLS;
            const int64_t b = (tile0 + (int64_t)gridDim.x * s) * C5_ROWS + row;
            float a0[16];
#pragma unroll
            for (int i = 0; i < 16; ++i) a0[i] = 0.0f;
            if (b < P.B) {
                a0[0] = __ldg(P.grad_sigma + b) * fminf(fmaxf(__ldg(P.sigma + b), e_lo), e_hi);  // trunc_exp backward (activation.py:13-17)
                if (P.grad_geo) {
#pragma unroll
                    for (int i = 0; i < 15; ++i) a0[1 + i] = __ldg(P.grad_geo + b * 15 + i);
                }
            }
            // the slot is free once the gather warps have taken the previous tile's feature gradients
            if (it > 0) tc5::mbar_wait_wd(tc5::smem_u32(&S.bars[s].t_free), (uint32_t)((it - 1) & 1));
            tc5::fence_after_sync();
            c5_put16(lane_addr, 0, a0);
            c5_chain_arrive(tc5::smem_u32(&S.bars[s].a_ready));
        }
        // ---- dH .* relu'
#pragma unroll 1
        for (int s = 0; s < n_act; ++s) {
            const uint32_t lane_addr = lane_base + s * C5_SLOT_COLS;
            const int64_t b = (tile0 + (int64_t)gridDim.x * s) * C5_ROWS + row;
            const uint64_t m = b < P.B ? __ldg(P.masks + b) : 0ull;
            tc5::mbar_wait_wd(tc5::smem_u32(&S.bars[s].d_full), (uint32_t)(it & 1));
            tc5::fence_after_sync();
            c5_mask_to_a(lane_addr, m);
            c5_chain_arrive(tc5::smem_u32(&S.bars[s].a_ready));
        }
    }
}

__device__ __forceinline__ void d5_backward_mma(D5Smem &S, const D5BwdParams &P, uint32_t tmem, int s) {
    const int64_t ntiles = (P.B + C5_ROWS - 1) / C5_ROWS;
    const uint32_t t = tmem + s * C5_SLOT_COLS;
    const uint32_t bar_d = tc5::smem_u32(&S.bars[s].d_full), bar_a = tc5::smem_u32(&S.bars[s].a_ready),
                   bar_g = tc5::smem_u32(&S.bars[s].g_full);
    const uint32_t s1h = tc5::smem_u32(S.w.s1[0]), s1l = tc5::smem_u32(S.w.s1[1]), s0h = tc5::smem_u32(S.w.s0[0]),
                   s0l = tc5::smem_u32(S.w.s0[1]);
    for (int64_t it = 0;; ++it) {
        const int64_t tile = blockIdx.x + (int64_t)gridDim.x * (2 * it + s);
        if (tile >= ntiles) break;
        const uint32_t pa = (uint32_t)(2 * it);
        tc5::mbar_wait_wd(bar_a, (pa + 0u) & 1u);
        tc5::fence_after_sync();
        c5_issue_layer<64, 16>(t + C5_D, t + C5_HHI, t + C5_HLO, s1h, s1l);   // dH = dO W_s1
        tc5::mma_commit(bar_d);
        tc5::mbar_wait_wd(bar_a, (pa + 1u) & 1u);
        tc5::fence_after_sync();
        c5_issue_layer<32, 64>(t + C5_D, t + C5_HHI, t + C5_HLO, s0h, s0l);   // dFeat = dHm W_s0
        tc5::mma_commit(bar_g);
    }
}

template <bool SAFE>
__device__ __forceinline__ void d5_backward_gather(D5Smem &S, const D5BwdParams &P, uint32_t tmem, int warp, int lane) {
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
k_dfield5_backward(D5BwdParams P, const int32_t *__restrict__ offsets, float S_, uint32_t H, uint32_t gridtype,
                   const float *__restrict__ W0, const float *__restrict__ W1) {
    extern __shared__ __align__(128) unsigned char d5_smem_raw[];
    D5Smem &S = *reinterpret_cast<D5Smem *>(d5_smem_raw);
    const int tid = threadIdx.x, warp = __shfl_sync(0xffffffffu, tid >> 5, 0), lane = tid & 31;
    fq_stage_levels(S.lv, &S.all_safe, offsets, S_, H, gridtype, P.align_corners);
    c5_stage_tile(S.w.s1[0], S.w.s1[1], 64, 16, [&](int n, int k) { return W1[k * FQ_HID + n]; });
    c5_stage_tile(S.w.s0[0], S.w.s0[1], 32, 64, [&](int n, int k) { return W0[k * FQ_F + n]; });
    if (tid == 0) d5_init_bars(S.bars);
    if (warp == C5_MMA_WARP) tc5::tmem_alloc(tc5::smem_u32(&S.tmem), 512);
    tc5::fence_async_smem();
    tc5::fence_before_sync();
    __syncthreads();
    tc5::fence_after_sync();
    const uint32_t tmem = __shfl_sync(0xffffffffu, S.tmem, 0);
    if (warp < C5_GATHER_WARPS) {
        if (S.all_safe) d5_backward_gather<true>(S, P, tmem, warp, lane);
        else d5_backward_gather<false>(S, P, tmem, warp, lane);
    } else if (warp < C5_MMA_WARP) {
        d5_backward_chain(S, P, tmem, warp, lane);
    } else {
        if (lane == 0) d5_backward_mma(S, P, tmem, warp - C5_MMA_WARP);
        __syncwarp();
    }
    tc5::fence_before_sync();
    __syncthreads();
    if (warp == C5_MMA_WARP) tc5::tmem_dealloc(tmem, 512);
}

// ------------------------------------------------------------------------------------------------ value + gradient
// sigma AND u * dsigma/dx in ONE table pass (the planner's collision cost, Jacobian rows, mean(sigma^2): fieldvg.cu states the
// contract).  sigma_i depends on x_i only, so dsigma/dx = dFeat . J with J = d(features)/dx (32 x 3 per point) known from the
// very corners the forward gathers: the gather thread writes J to shared memory next to the features, the tensor cores run
// L0, L1 and — from the chain's (relu' .* W1[0,:]) * u sigma' — the feature-gradient product, and the chain thread of the
// row contracts those 32 feature gradients with its row of J; nothing is gathered twice.  J is one 48 KB tile per slot: the
// next gather of the slot waits for the chain of the previous tile (the other slot's gathers fill the L1 pipe meanwhile).  A
// ring of three J tiles removes that wait but costs the gathers their L1: measured on the forward kernel, the same launch
// with 120 / 150 / 180 / 205 KB of shared memory runs 1.338 / 1.356 / 1.457 / 2.625 ms — outstanding gather misses live in
// L1 lines, so the shared-memory carve-out is part of the kernel's roofline (the ring version ran 2.88 ms at 203 KB).
struct __align__(128) D5VgSmem {
    D5Weights w;             // s0 [64 x 32], s1 [16 x 64]  (forward tiles)
    float b0[2][32 * 64];    // dFeat = dHm W_s0: the tile is W_s0^T [32 x 64]
    float w1row[FQ_HID];     // W_s1[0, :] = d(log sigma)/d(hidden)
    FqLevel lv[FQ_L];
    C5Bars bars[2];
    uint64_t j_free[2];      // 128 chain threads: the slot's J rows have been contracted
    uint32_t tmem;
    int all_safe;
    __align__(128) float stage[4][2][32 * 15];   // geo_feat rows of a chain warp (bulk store)
    __align__(128) float gstage[4][2][32 * 3];   // grad_x rows of a chain warp (bulk store)
    __align__(128) float jac[2][FQ_F * 3][C5_ROWS];
};
struct D5VgParams {
    const float *x, *table;
    float bound;
    int align_corners;
    const float *coef_a, *coef_b;
    float a0, b0;
    float *sigma, *geo, *grad_x;
    unsigned long long *masks;   // optional: the ReLU sign bits k_dfield5_backward reads (same layout as k_dfield5_forward's)
    int64_t B;
};

// levels 2*lp, 2*lp+1 of one sample: the 4 interpolated features (c5_encode_pair) and their 12 partial derivatives
// jc[(2 h + c) * 3 + axis] (c5_encode_pair_backward's per-axis sums, same expressions) from one set of 16 gathers
template <bool SAFE>
__device__ __forceinline__ void d5_encode_pair_jac(const float *__restrict__ table, const FqLevel *__restrict__ lvs, int lp,
                                                   const float (&x01)[3], bool valid, int align_corners, float (&feat)[4],
                                                   float (&jc)[12]) {
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
            jc[(2 * h) * 3 + gd] = valid ? a0 : 0.0f;
            jc[(2 * h + 1) * 3 + gd] = valid ? a1 : 0.0f;
        }
    }
}

// All eight gather warps work on ONE tile: the two threads of a row (warps q and q + 4) take the even / odd level pairs, and
// consecutive tiles alternate between the two TMEM slots / J tiles.  The chain of tile n then runs during the gather of tile
// n + 1, and tile n + 2 finds its slot free.  (With one tile per four-warp group, as in the forward kernel, both groups waited
// for their J tile at the same time: D5_STATS showed 27 % of the gather warps' cycles in that wait, 1.86 ms per launch.)
template <bool SAFE>
__device__ __forceinline__ void d5_vg_gather(D5VgSmem &S, const D5VgParams &P, uint32_t tmem, int warp, int lane) {
    const int g = warp >> 2, q = warp & 3, row = 32 * q + lane;
    const int64_t ntiles = (P.B + C5_ROWS - 1) / C5_ROWS;
#if D5_STATS
    long long w_fe = 0, w_jf = 0;
    const long long t_start = clock64();
#endif
    for (int64_t n = 0;; ++n) {
        const int64_t tile = blockIdx.x + (int64_t)gridDim.x * n;
        if (tile >= ntiles) {
#if D5_STATS
            if (blockIdx.x == 0 && lane == 0 && q == 0)
                printf("d5vg gather half %d: tiles %lld total %lld cyc, wait f_empty %lld, wait j_free %lld\n", g, (long long)n,
                       clock64() - t_start, w_fe, w_jf);
#endif
            break;
        }
        const int s = (int)(n & 1);
        const int64_t it = n >> 1;
        const uint32_t lane_addr = tmem + s * C5_SLOT_COLS + ((uint32_t)(32 * q) << 16);
        float *jrow = &S.jac[s][0][row];
        const int64_t b = tile * C5_ROWS + row;
        float x01[3];
        const bool valid = fq_load_point(P.x, b, P.B, P.bound, x01);
#pragma unroll 1
        for (int k = 0; k < 4; ++k) {
            const int lp = 2 * k + g;
            float feat[4], jc[12];
            d5_encode_pair_jac<SAFE>(P.table, S.lv, lp, x01, valid, P.align_corners, feat, jc);
            if (k == 0 && it > 0) {
                D5_TIMED(w_fe, tc5::mbar_wait_wd(tc5::smem_u32(&S.bars[s].f_empty), (uint32_t)((it - 1) & 1)));
                D5_TIMED(w_jf, tc5::mbar_wait_wd(tc5::smem_u32(&S.j_free[s]), (uint32_t)((it - 1) & 1)));
                tc5::fence_after_sync();
            }
            c5_put_features(lane_addr, lp, feat);
#pragma unroll
            for (int i = 0; i < 12; ++i) jrow[(12 * lp + i) * C5_ROWS] = jc[i];
        }
        tc5::tmem_st_wait();
        tc5::fence_before_sync();
        tc5::mbar_arrive(tc5::smem_u32(&S.bars[s].f_full));
    }
}

// A operand of the feature-gradient product: dH_j = c * W_s1[0, j] where hidden unit j was positive, else 0
static __device__ __noinline__ void d5_vg_put_dh(uint32_t lane_addr, uint64_t m, float c, const float *w1row) {
#pragma unroll 1
    for (int ch = 0; ch < 4; ++ch) {
        float v[16];
        const int bits = (int)(uint32_t)(m >> (16 * ch));
#pragma unroll
        for (int i = 0; i < 16; ++i) {
            int keep;  // 0 or -1: sign-extended bit i
            asm("bfe.s32 %0, %1, %2, 1;" : "=r"(keep) : "r"(bits), "r"(i));
            v[i] = __uint_as_float(__float_as_uint(c * w1row[16 * ch + i]) & (uint32_t)keep);
        }
        c5_put16(lane_addr, 16 * ch, v);
    }
}

// chain warps: the three epilogues of a tile, tile after tile (d_full completes three times per tile, a_ready twice)
__device__ __forceinline__ void d5_vg_chain(D5VgSmem &S, const D5VgParams &P, uint32_t tmem, int warp, int lane) {
    const int q = warp & 3, row = 32 * q + lane;
    const int64_t ntiles = (P.B + C5_ROWS - 1) / C5_ROWS;
    const float e_lo = expf(-15.0f), e_hi = expf(15.0f), den = 2.0f * P.bound;
    const uint32_t lane_base = tmem + ((uint32_t)(32 * q) << 16);
    const bool geo_bulk = P.geo && (reinterpret_cast<uintptr_t>(P.geo) & 15) == 0;
    const bool gx_bulk = (reinterpret_cast<uintptr_t>(P.grad_x) & 15) == 0;
#if D5_STATS
    long long c_ph[3] = {0, 0, 0}, c_w[3] = {0, 0, 0};
    const long long t_start = clock64();
#endif
    for (int64_t n = 0;; ++n) {
        const int64_t tile = blockIdx.x + (int64_t)gridDim.x * n;
        if (tile >= ntiles) break;
        const int s = (int)(n & 1);
        const uint32_t par = (uint32_t)(3 * (n >> 1));
        const uint32_t lane_addr = lane_base + s * C5_SLOT_COLS, bar_d = tc5::smem_u32(&S.bars[s].d_full),
                       bar_a = tc5::smem_u32(&S.bars[s].a_ready);
        const int64_t b = tile * C5_ROWS + row, b0 = b - lane;
        const bool inb = b < P.B, full = b0 + 32 <= P.B;
        // ---- hidden layer: ReLU -> A operand of the output layer; the sign bits stay in a register
        D5_TIMED(c_w[0], tc5::mbar_wait_wd(bar_d, (par + 0u) & 1u));
        tc5::fence_after_sync();
        uint64_t m;
        D5_TIMED(c_ph[0], m = c5_relu_to_a<true>(lane_addr); c5_chain_arrive(bar_a));
        if (P.masks && inb) P.masks[b] = m;
        // ---- output layer: sigma, geo_feat; A operand of the feature-gradient product
        D5_TIMED(c_w[1], tc5::mbar_wait_wd(bar_d, (par + 1u) & 1u));
        tc5::fence_after_sync();
#if D5_STATS
        long long t_ph = clock64();
#endif
        float o[16];
        c5_get16(lane_addr, 0, o);
        const float sg = expf(o[0]);  // trunc_exp forward (activation.py:9-11)
        const float ua = (inb && P.coef_a) ? __ldg(P.coef_a + b) : P.a0, ub = (inb && P.coef_b) ? __ldg(P.coef_b + b) : P.b0;
        const float c = inb ? fmaf(ub, sg, ua) * fminf(fmaxf(sg, e_lo), e_hi) : 0.0f;  // u * trunc_exp backward (:13-17)
        d5_vg_put_dh(lane_addr, m, c, S.w1row);
        c5_chain_arrive(bar_a);
        if (inb) P.sigma[b] = sg;
        if (geo_bulk && full) {
            float *st = S.stage[q][s];
            if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");  // the buffer's last copy (two tiles ago)
            __syncwarp();
#pragma unroll
            for (int i = 0; i < 15; ++i) st[lane * 15 + i] = o[1 + i];
            tc5::fence_async_smem();
            __syncwarp();
            if (lane == 0) d5_bulk_store(P.geo + b0 * 15, st, 32 * 15 * 4);
        } else if (P.geo && inb) {
#pragma unroll
            for (int i = 0; i < 15; ++i) P.geo[b * 15 + i] = o[1 + i];
        }
#if D5_STATS
        c_ph[1] += clock64() - t_ph;
#endif
        // ---- feature gradients . J -> u dsigma/dx  (the order of c5_encode_pair_backward's sums)
        D5_TIMED(c_w[2], tc5::mbar_wait_wd(bar_d, (par + 2u) & 1u));
        tc5::fence_after_sync();
#if D5_STATS
        t_ph = clock64();
#endif
        const float *jrow = &S.jac[s][0][row];
        float gx[3] = {0.0f, 0.0f, 0.0f};
#pragma unroll 1
        for (int h = 0; h < 2; ++h) {
            uint32_t r[16];
            tc5::tmem_ld16_nowait(lane_addr + C5_D + 16 * h, r);
            tc5::tmem_ld_wait();
#pragma unroll
            for (int f = 0; f < 16; ++f) {
#pragma unroll
                for (int gd = 0; gd < 3; ++gd)
                    gx[gd] = fmaf(__uint_as_float(r[f]), jrow[(48 * h + 3 * f + gd) * C5_ROWS], gx[gd]);
            }
        }
        tc5::fence_before_sync();
        tc5::mbar_arrive(tc5::smem_u32(&S.bars[s].t_free));
        tc5::mbar_arrive(tc5::smem_u32(&S.j_free[s]));
        if (gx_bulk && full) {
            float *st = S.gstage[q][s];
            if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
            __syncwarp();
#pragma unroll
            for (int a = 0; a < 3; ++a) st[lane * 3 + a] = __fdiv_rn(gx[a], den);
            tc5::fence_async_smem();
            __syncwarp();
            if (lane == 0) d5_bulk_store(P.grad_x + b0 * 3, st, 32 * 3 * 4);
        } else if (inb) {
#pragma unroll
            for (int a = 0; a < 3; ++a) P.grad_x[b * 3 + a] = __fdiv_rn(gx[a], den);
        }
#if D5_STATS
        c_ph[2] += clock64() - t_ph;
#endif
    }
#if D5_STATS
    if (blockIdx.x == 0 && lane == 0 && q == 0)
        printf("d5vg chain: total %lld cyc, phase A %lld (wait %lld), B %lld (wait %lld), C %lld (wait %lld)\n", clock64() - t_start,
               c_ph[0], c_w[0], c_ph[1], c_w[1], c_ph[2], c_w[2]);
#endif
    if (lane == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}

__device__ __forceinline__ void d5_vg_mma(D5VgSmem &S, const D5VgParams &P, uint32_t tmem, int s) {
    const int64_t ntiles = (P.B + C5_ROWS - 1) / C5_ROWS;
    const uint32_t t = tmem + s * C5_SLOT_COLS;
    const uint32_t bar_ff = tc5::smem_u32(&S.bars[s].f_full), bar_fe = tc5::smem_u32(&S.bars[s].f_empty),
                   bar_d = tc5::smem_u32(&S.bars[s].d_full), bar_a = tc5::smem_u32(&S.bars[s].a_ready),
                   bar_t = tc5::smem_u32(&S.bars[s].t_free);
    const uint32_t s0h = tc5::smem_u32(S.w.s0[0]), s0l = tc5::smem_u32(S.w.s0[1]), s1h = tc5::smem_u32(S.w.s1[0]),
                   s1l = tc5::smem_u32(S.w.s1[1]), b0h = tc5::smem_u32(S.b0[0]), b0l = tc5::smem_u32(S.b0[1]);
#if D5_STATS
    long long w_ff = 0, w_tf = 0, w_a0 = 0, w_a1 = 0;
    const long long t_start = clock64();
#endif
    for (int64_t it = 0;; ++it) {
        const int64_t tile = blockIdx.x + (int64_t)gridDim.x * (2 * it + s);
        if (tile >= ntiles) {
#if D5_STATS
            if (blockIdx.x == 0)
                printf("d5vg mma slot %d: total %lld cyc, wait f_full %lld, t_free %lld, a_ready(L1) %lld, a_ready(dF) %lld\n", s,
                       clock64() - t_start, w_ff, w_tf, w_a0, w_a1);
#endif
            break;
        }
        const uint32_t pa = (uint32_t)(2 * it);
        D5_TIMED(w_ff, tc5::mbar_wait_wd(bar_ff, (uint32_t)(it & 1)));
        if (it > 0) D5_TIMED(w_tf, tc5::mbar_wait_wd(bar_t, (uint32_t)((it - 1) & 1)));
        tc5::fence_after_sync();
        c5_issue_layer<64, 32>(t + C5_D, t + C5_FHI, t + C5_FLO, s0h, s0l);   // features -> 64
        tc5::mma_commit(bar_fe);
        tc5::mma_commit(bar_d);
        D5_TIMED(w_a0, tc5::mbar_wait_wd(bar_a, (pa + 0u) & 1u));
        tc5::fence_after_sync();
        c5_issue_layer<16, 64>(t + C5_D, t + C5_HHI, t + C5_HLO, s1h, s1l);   // 64 -> 16
        tc5::mma_commit(bar_d);
        D5_TIMED(w_a1, tc5::mbar_wait_wd(bar_a, (pa + 1u) & 1u));
        tc5::fence_after_sync();
        c5_issue_layer<32, 64>(t + C5_D, t + C5_HHI, t + C5_HLO, b0h, b0l);   // dFeat = dHm W_s0
        tc5::mma_commit(bar_d);
    }
}

__global__ void __launch_bounds__(C5_THREADS, 1)
k_dfield5_value_grad(D5VgParams P, const int32_t *__restrict__ offsets, float S_, uint32_t H, uint32_t gridtype,
                     const float *__restrict__ W0, const float *__restrict__ W1) {
    extern __shared__ __align__(128) unsigned char d5_smem_raw[];
    D5VgSmem &S = *reinterpret_cast<D5VgSmem *>(d5_smem_raw);
    const int tid = threadIdx.x, warp = __shfl_sync(0xffffffffu, tid >> 5, 0), lane = tid & 31;
    fq_stage_levels(S.lv, &S.all_safe, offsets, S_, H, gridtype, P.align_corners);
    c5_stage_tile(S.w.s0[0], S.w.s0[1], 64, 32, [&](int n, int k) { return W0[n * FQ_F + k]; });
    c5_stage_tile(S.w.s1[0], S.w.s1[1], 16, 64, [&](int n, int k) { return W1[n * FQ_HID + k]; });
    c5_stage_tile(S.b0[0], S.b0[1], 32, 64, [&](int n, int k) { return W0[k * FQ_F + n]; });
    if (tid < FQ_HID) S.w1row[tid] = W1[tid];
    if (tid == 0) {
        d5_init_bars(S.bars, 256);   // every gather thread arrives on the tile's f_full
        for (int j = 0; j < 2; ++j) tc5::mbar_init(tc5::smem_u32(&S.j_free[j]), 128);
        tc5::fence_mbar_init();
    }
    if (warp == C5_MMA_WARP) tc5::tmem_alloc(tc5::smem_u32(&S.tmem), 512);
    tc5::fence_async_smem();
    tc5::fence_before_sync();
    __syncthreads();
    tc5::fence_after_sync();
    const uint32_t tmem = __shfl_sync(0xffffffffu, S.tmem, 0);
    if (warp < C5_GATHER_WARPS) {
        if (S.all_safe) d5_vg_gather<true>(S, P, tmem, warp, lane);
        else d5_vg_gather<false>(S, P, tmem, warp, lane);
    } else if (warp < C5_MMA_WARP) {
        d5_vg_chain(S, P, tmem, warp, lane);
    } else {
        if (lane == 0) d5_vg_mma(S, P, tmem, warp - C5_MMA_WARP);
        __syncwarp();
    }
    tc5::fence_before_sync();
    __syncthreads();
    if (warp == C5_MMA_WARP) tc5::tmem_dealloc(tmem, 512);
}

// ------------------------------------------------------------------------------------------------ launchers (field.cu)
// one CTA per SM: a CTA owns all 512 TMEM columns (the shared-memory request keeps a second CTA off the SM)
static int d5_smem_bytes() {
    const char *e = getenv("NGP_D5_SMEM_KB");  // experiment: L1 capacity left to the gathers
    if (e) return atoi(e) * 1024;
    return (int)(sizeof(D5Smem) > 120 * 1024 ? sizeof(D5Smem) : 120 * 1024);
}
static int d5_grid(int64_t B) {
    int dev = 0, sms = NGP_SM_COUNT;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const int64_t tiles = ngp_div_up(B, C5_ROWS);
    return (int)(tiles < sms ? tiles : sms);
}

int d5_field_query_forward(const float *x, float bound, const float *embeddings, const int32_t *offsets, float S, uint32_t H,
                           uint32_t gridtype, int align_corners, const float *W0, const float *W1, float *sigma, float *geo,
                           uint64_t *relu_mask, int64_t B, cudaStream_t stream) {
    const int smem = d5_smem_bytes();
    if (cudaFuncSetAttribute(k_dfield5_forward, cudaFuncAttributeMaxDynamicSharedMemorySize, smem) != cudaSuccess) {
        ngp_set_error("ngp_field_query_forward: cannot reserve %d B of shared memory", smem);
        return NGP_ERR_CUDA;
    }
    D5FwdParams P;
    P.x = x; P.table = embeddings; P.bound = bound; P.align_corners = align_corners;
    P.sigma = sigma; P.geo = geo; P.masks = (unsigned long long *)relu_mask; P.B = B;
    k_dfield5_forward<<<d5_grid(B), C5_THREADS, smem, stream>>>(P, offsets, S, H, gridtype, W0, W1);
    NGP_CHECK_LAUNCH("ngp_field_query_forward (tcgen05)");
    return NGP_OK;
}

int d5_field_query_backward(const float *x, float bound, const float *embeddings, const int32_t *offsets, float S, uint32_t H,
                            uint32_t gridtype, int align_corners, const float *W0, const float *W1, const float *sigma,
                            const uint64_t *relu_mask, const float *grad_sigma, const float *grad_geo, float *grad_x,
                            int64_t B, cudaStream_t stream) {
    const int smem = d5_smem_bytes();
    if (cudaFuncSetAttribute(k_dfield5_backward, cudaFuncAttributeMaxDynamicSharedMemorySize, smem) != cudaSuccess) {
        ngp_set_error("ngp_field_query_backward: cannot reserve %d B of shared memory", smem);
        return NGP_ERR_CUDA;
    }
    D5BwdParams P;
    P.x = x; P.table = embeddings; P.bound = bound; P.align_corners = align_corners;
    P.sigma = sigma; P.masks = (const unsigned long long *)relu_mask; P.grad_sigma = grad_sigma; P.grad_geo = grad_geo;
    P.grad_x = grad_x; P.B = B;
    k_dfield5_backward<<<d5_grid(B), C5_THREADS, smem, stream>>>(P, offsets, S, H, gridtype, W0, W1);
    NGP_CHECK_LAUNCH("ngp_field_query_backward (tcgen05)");
    return NGP_OK;
}

int d5_field_query_value_grad(const float *x, float bound, const float *embeddings, const int32_t *offsets, float S, uint32_t H,
                              uint32_t gridtype, int align_corners, const float *W0, const float *W1, const float *coef_a,
                              const float *coef_b, float a0, float b0, float *sigma, float *geo, float *grad_x,
                              uint64_t *relu_mask, int64_t B, cudaStream_t stream) {
    const int smem = (int)sizeof(D5VgSmem);
    if (cudaFuncSetAttribute(k_dfield5_value_grad, cudaFuncAttributeMaxDynamicSharedMemorySize, smem) != cudaSuccess) {
        ngp_set_error("ngp_field_query_value_grad: cannot reserve %d B of shared memory", smem);
        return NGP_ERR_CUDA;
    }
    D5VgParams P;
    P.x = x; P.table = embeddings; P.bound = bound; P.align_corners = align_corners;
    P.coef_a = coef_a; P.coef_b = coef_b; P.a0 = a0; P.b0 = b0;
    P.sigma = sigma; P.geo = geo; P.grad_x = grad_x; P.masks = (unsigned long long *)relu_mask; P.B = B;
    k_dfield5_value_grad<<<d5_grid(B), C5_THREADS, smem, stream>>>(P, offsets, S, H, gridtype, W0, W1);
    NGP_CHECK_LAUNCH("ngp_field_query_value_grad");
    return NGP_OK;
}
