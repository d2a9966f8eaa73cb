// ffmlp5.cu — warp-specialised forward / inference kernel of the fully fused 64-wide fp16 MLP (the reference's
// ffmlp_forward / ffmlp_inference, ffmlp/src/ffmlp.cu:331-446, host :525-700; Python ffmlp/ffmlp.py:15-77).
// Same semantics as ffmlp.cu's first kernel (flat fp16 weights [64 x in][64 x 64]*(n-1)[16 x 64], ReLU or no hidden
// activation, activations rounded to fp16 after every layer, forward_buffer [n, B, 64] when training, outputs [B, 16]),
// different machine mapping.  That kernel ran one 128-row tile per 128-thread CTA with a CTA-wide barrier around every
// layer: measured 0.52 ms per 4,194,304 rows = 0.12 of the HBM copy rate, bound by the latency of its serial
// load -> MMA -> epilogue chain, not by memory.  Here ONE persistent CTA per SM keeps up to 8 tiles in flight:
//
//   loader warps (4)    thread = tile row: the row's input chunks go global -> shared with cp.async (16 B each, no
//                       registers, zero fill past B) into the chunk-major K-major A-operand tile of a slot; the copies'
//                       completion arrives on the slot's mbarrier (cp.async.mbarrier.arrive.noinc), so the loader runs
//                       ahead over every free slot and many rows' worth of loads are in flight per SM;
//   MMA threads (4)     each walks its slots layer by layer: tcgen05.mma kind::f16 (A and the weight tiles in shared memory,
//                       fp32 accumulator of the slot in TMEM, 64 columns per slot), tcgen05.commit onto the slot's mbarrier;
//   epilogue warps (8)  two groups of four (a warp reads the 32 TMEM lanes of its quadrant): tcgen05.ld -> activation ->
//                       fp16 -> the slot's hidden tile in shared memory (the next layer's A operand) and, when training,
//                       forward_buffer; after the output layer the 16 columns go to `outputs`.
//
// While the epilogue of one slot runs, the MMA thread issues the same layer of the other slots, and the loaders refill the
// input tile of a slot as soon as its first layer has been read (x_empty), so layers, epilogues and HBM loads overlap
// across tiles instead of queueing behind a barrier.
#include "cfield5.cuh"   // common.cuh, tc5.cuh, the hash-grid gather helpers of the field kernels (fused encode -> FFMLP below)

#define F5_ROWS 128
#define F5_HID 64
#define F5_OUT 16
#define F5_ACT_RELU 0
#define F5_LOAD_WARPS 4
#ifndef F5_EPI_GROUPS
#define F5_EPI_GROUPS 2
#endif
#ifndef F5_MMA_WARPS
#define F5_MMA_WARPS 4          // one issuing thread per warp, slots dealt round-robin (see f5_mma)
#endif
#define F5_MMA_WARP F5_LOAD_WARPS
#define F5_EPI_WARP0 (F5_MMA_WARP + F5_MMA_WARPS)
#define F5_THREADS ((F5_EPI_WARP0 + 4 * F5_EPI_GROUPS) * 32)   // 512
#define F5_MAX_SLOTS 8

struct F5Bars {
    uint64_t x_full;    // 128 loader threads (cp.async completion): the slot's input tile is in shared memory
    uint64_t x_empty;   // tcgen05.commit: the first layer has read the input tile
    uint64_t d_full;    // tcgen05.commit: a layer's accumulator is complete
    uint64_t a_ready;   // 128 epilogue threads: accumulator read, the next A operand (or nothing, after the output) written
};
struct F5Shared {
    F5Bars bars[F5_MAX_SLOTS];
    uint32_t tmem;
};

struct F5Params {
    const __half *x, *w;
    __half *fwd_buf, *out;
    int64_t B;
    int in_dim, n_layers, act, slots;
};

__device__ __forceinline__ void f5_cp_async16(uint32_t dst, const void *src, bool valid) {
    const uint32_t n = valid ? 16u : 0u;  // src-size 0: the 16 bytes are zero-filled
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(n) : "memory");
}
// 32 contiguous bytes per thread in ONE store (st.global.v8.b32, sm_100): a full sector per lane.  Two 16-byte stores reach
// L2 as two partial-sector writes each.
__device__ __forceinline__ void f5_st32(void *dst, const uint4 &a, const uint4 &b) {
    asm volatile("st.global.v8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"l"(dst), "r"(a.x), "r"(a.y), "r"(a.z), "r"(a.w),
                 "r"(b.x), "r"(b.y), "r"(b.z), "r"(b.w)
                 : "memory");
}
__device__ __forceinline__ void f5_cp_async_arrive(uint32_t bar) {
    asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];" ::"r"(bar) : "memory");
}

// stage a row-major [N, K] fp16 weight block as a K-major chunk-major tile (rows = N)
__device__ __forceinline__ void f5_stage_weights(unsigned char *dst, const __half *__restrict__ w, int N, int K) {
    const int kcs = K >> 3;
    for (int i = threadIdx.x; i < N * kcs; i += blockDim.x) {
        const int n = i / kcs, kc = i - n * kcs;
        *reinterpret_cast<uint4 *>(dst + tc5::chunk_off(N, n, kc)) = __ldg(reinterpret_cast<const uint4 *>(w + (size_t)n * K + kc * 8));
    }
}

__device__ __forceinline__ void f5_issue_layer(uint32_t d_tmem, uint32_t a_addr, uint32_t w_addr, int N, int K) {
    const uint32_t idesc = tc5::idesc_f16(F5_ROWS, N, tc5::MAJOR_K, tc5::MAJOR_K);
    for (int ks = 0; ks < (K >> 4); ++ks) {
        const uint64_t ad = tc5::smem_desc(a_addr + ks * 2 * F5_ROWS * 16, F5_ROWS * 16, 128);
        const uint64_t bd = tc5::smem_desc(w_addr + ks * 2 * N * 16, N * 16, 128);
        tc5::mma_f16(d_tmem, ad, bd, idesc, ks > 0);
    }
}

// tiles of this CTA: number `n` (0, 1, ...) is global tile blockIdx.x + gridDim.x * n; round r holds n = r * slots + s
__device__ __forceinline__ int64_t f5_tile(const F5Params &P, int64_t r, int s) {
    return blockIdx.x + (int64_t)gridDim.x * (r * P.slots + s);
}

// ------------------------------------------------------------------------------------------------ roles
__device__ __forceinline__ void f5_loader(const F5Params &P, F5Shared *sh, unsigned char *Xs, int tid) {
    const int64_t ntiles = (P.B + F5_ROWS - 1) / F5_ROWS;
    const int kcs = P.in_dim >> 3;
    const uint32_t x_bytes = (uint32_t)F5_ROWS * P.in_dim * 2;
    for (int64_t r = 0;; ++r) {
        if (f5_tile(P, r, 0) >= ntiles) break;
        for (int s = 0; s < P.slots; ++s) {
            const int64_t tile = f5_tile(P, r, s);
            if (tile >= ntiles) break;
            if (r > 0) tc5::mbar_wait_wd(tc5::smem_u32(&sh->bars[s].x_empty), (uint32_t)((r - 1) & 1));
            const uint32_t dst = tc5::smem_u32(Xs) + s * x_bytes;
            // the tile's 16-byte chunks in memory order, 128 per pass: a warp's 32 copies read 512 contiguous bytes
            for (int i = 0; i < kcs; ++i) {
                const int f = i * F5_ROWS + tid, trow = f / kcs, kc = f - trow * kcs;
                const int64_t row = tile * F5_ROWS + trow;
                const bool inb = row < P.B;
                f5_cp_async16(dst + tc5::chunk_off(F5_ROWS, trow, kc), P.x + (inb ? row : 0) * P.in_dim + kc * 8, inb);
            }
            f5_cp_async_arrive(tc5::smem_u32(&sh->bars[s].x_full));
        }
    }
    asm volatile("cp.async.wait_all;" ::: "memory");
}

// One thread spends ~90 cycles per tcgen05.mma on descriptor set-up (uniform-register moves), ~900 per (slot, layer) with
// the waits and commits, while the tensor core needs ~130: with a single issuing thread the kernel was bound by that thread
// (0.31 ms per 4,194,304 rows whether 2, 3 or 4 epilogue groups ran).  F5_MMA_WARPS threads in different warps, thread k
// owning the slots s % F5_MMA_WARPS == k, issue side by side; each thread's tcgen05.commit covers its own MMAs.
__device__ __forceinline__ void f5_mma(const F5Params &P, F5Shared *sh, uint32_t tmem, uint32_t xs_a, uint32_t hs_a,
                                       uint32_t win_a, uint32_t wh_a, uint32_t wout_a, int k, int n_mma = F5_MMA_WARPS) {
    const int64_t ntiles = (P.B + F5_ROWS - 1) / F5_ROWS;
    const int n = P.n_layers;
    const uint32_t x_bytes = (uint32_t)F5_ROWS * P.in_dim * 2, h_bytes = F5_ROWS * F5_HID * 2;
    for (int64_t r = 0;; ++r) {
        if (f5_tile(P, r, 0) >= ntiles) break;
        for (int l = 0; l <= n; ++l) {
            for (int s = k; s < P.slots; s += n_mma) {
                if (f5_tile(P, r, s) >= ntiles) break;
                const uint32_t bar_a = tc5::smem_u32(&sh->bars[s].a_ready), bar_d = tc5::smem_u32(&sh->bars[s].d_full);
                const uint32_t acc = tmem + s * F5_HID;
                if (l == 0) {
                    tc5::mbar_wait_wd(tc5::smem_u32(&sh->bars[s].x_full), (uint32_t)(r & 1));
                    if (r > 0) tc5::mbar_wait_wd(bar_a, (uint32_t)((r * (n + 1) - 1) & 1));  // the slot's last output has been read
                    tc5::fence_async_smem();  // cp.async wrote the tile through the generic proxy; the MMA reads it through the async one
                    tc5::fence_after_sync();
                    f5_issue_layer(acc, xs_a + s * x_bytes, win_a, F5_HID, P.in_dim);
                    tc5::mma_commit(tc5::smem_u32(&sh->bars[s].x_empty));
                } else {
                    tc5::mbar_wait_wd(bar_a, (uint32_t)((r * (n + 1) + l - 1) & 1));
                    tc5::fence_after_sync();
                    if (l < n) f5_issue_layer(acc, hs_a + s * h_bytes, wh_a + (l - 1) * F5_HID * F5_HID * 2, F5_HID, F5_HID);
                    else f5_issue_layer(acc, hs_a + s * h_bytes, wout_a, F5_OUT, F5_HID);
                }
                tc5::mma_commit(bar_d);
            }
        }
    }
}

// epilogue group g of n_groups serves the slots s = g, g + n_groups, ...; `warp` only selects the TMEM quadrant (warp % 4)
__device__ __forceinline__ void f5_epilogue(const F5Params &P, F5Shared *sh, uint32_t tmem, unsigned char *Hs, int warp, int lane,
                                            int g, int n_groups = F5_EPI_GROUPS) {
    const int64_t ntiles = (P.B + F5_ROWS - 1) / F5_ROWS;
    const int n = P.n_layers;
    const int q = warp & 3, trow = 32 * q + lane;
    const uint32_t lane_base = tmem + ((uint32_t)(32 * q) << 16);
    for (int64_t r = 0;; ++r) {
        if (f5_tile(P, r, 0) >= ntiles) break;
        for (int l = 0; l <= n; ++l) {
            for (int s = g; s < P.slots; s += n_groups) {
                const int64_t tile = f5_tile(P, r, s);
                if (tile >= ntiles) break;
                const int64_t row = tile * F5_ROWS + trow;
                const bool inb = row < P.B;
                tc5::mbar_wait_wd(tc5::smem_u32(&sh->bars[s].d_full), (uint32_t)((r * (n + 1) + l) & 1));
                tc5::fence_after_sync();
                const uint32_t acc = lane_base + s * F5_HID;
                if (l < n) {
                    unsigned char *H = Hs + (size_t)s * F5_ROWS * F5_HID * 2;
                    uint32_t raw[4][16];
#pragma unroll
                    for (int c = 0; c < 4; ++c) tc5::tmem_ld16_nowait(acc + 16 * c, raw[c]);
                    tc5::tmem_ld_wait();
#pragma unroll
                    for (int c = 0; c < 4; ++c) {
                        float v[16];
#pragma unroll
                        for (int i = 0; i < 16; ++i) {
                            v[i] = __uint_as_float(raw[c][i]);
                            if (P.act == F5_ACT_RELU) v[i] = fmaxf(v[i], 0.0f);
                        }
                        const uint4 lo = tc5::pack8(v, 0), hi = tc5::pack8(v, 8);
                        *reinterpret_cast<uint4 *>(H + tc5::chunk_off(F5_ROWS, trow, 2 * c)) = lo;
                        *reinterpret_cast<uint4 *>(H + tc5::chunk_off(F5_ROWS, trow, 2 * c + 1)) = hi;
                        if (P.fwd_buf && inb) f5_st32(P.fwd_buf + ((size_t)l * P.B + row) * F5_HID + c * 16, lo, hi);
                    }
                    tc5::fence_async_smem();   // the next layer's MMA reads the tile through the async proxy
                } else {
                    float v[16];
                    tc5::tmem_ld16(acc, v);
                    if (inb) f5_st32(P.out + row * F5_OUT, tc5::pack8(v, 0), tc5::pack8(v, 8));
                }
                tc5::fence_before_sync();
                tc5::mbar_arrive(tc5::smem_u32(&sh->bars[s].a_ready));
            }
        }
    }
}

__global__ void __launch_bounds__(F5_THREADS, 1) k_ffmlp5_forward(F5Params P, uint32_t tmem_cols) {
    extern __shared__ __align__(1024) unsigned char f5_smem[];
    F5Shared *sh = reinterpret_cast<F5Shared *>(f5_smem);
    const int n = P.n_layers;
    unsigned char *Win = f5_smem + 1024;                                   // [64 x in]
    unsigned char *Wh = Win + F5_HID * P.in_dim * 2;                        // (n-1) x [64 x 64]
    unsigned char *Wout = Wh + (n - 1) * F5_HID * F5_HID * 2;               // [16 x 64]
    unsigned char *Xs = Wout + F5_OUT * F5_HID * 2;                         // slots x [128 x in]
    unsigned char *Hs = Xs + (size_t)P.slots * F5_ROWS * P.in_dim * 2;      // slots x [128 x 64]
    const int tid = threadIdx.x, warp = __shfl_sync(0xffffffffu, tid >> 5, 0), lane = tid & 31;
    if (tid == 0) {
        for (int s = 0; s < P.slots; ++s) {
            tc5::mbar_init(tc5::smem_u32(&sh->bars[s].x_full), F5_LOAD_WARPS * 32);
            tc5::mbar_init(tc5::smem_u32(&sh->bars[s].x_empty), 1);
            tc5::mbar_init(tc5::smem_u32(&sh->bars[s].d_full), 1);
            tc5::mbar_init(tc5::smem_u32(&sh->bars[s].a_ready), 128);
        }
        tc5::fence_mbar_init();
    }
    if (warp == F5_MMA_WARP) tc5::tmem_alloc(tc5::smem_u32(&sh->tmem), tmem_cols);
    f5_stage_weights(Win, P.w, F5_HID, P.in_dim);
    for (int l = 0; l < n - 1; ++l)
        f5_stage_weights(Wh + l * F5_HID * F5_HID * 2, P.w + F5_HID * P.in_dim + l * F5_HID * F5_HID, F5_HID, F5_HID);
    f5_stage_weights(Wout, P.w + F5_HID * P.in_dim + (n - 1) * F5_HID * F5_HID, F5_OUT, F5_HID);
    tc5::fence_async_smem();
    tc5::fence_before_sync();
    __syncthreads();
    tc5::fence_after_sync();
    const uint32_t tmem = __shfl_sync(0xffffffffu, sh->tmem, 0);
    if (warp < F5_LOAD_WARPS) {
        f5_loader(P, sh, Xs, tid);
    } else if (warp < F5_EPI_WARP0) {
        if (lane == 0)
            f5_mma(P, sh, tmem, tc5::smem_u32(Xs), tc5::smem_u32(Hs), tc5::smem_u32(Win), tc5::smem_u32(Wh), tc5::smem_u32(Wout),
                   warp - F5_MMA_WARP);
        __syncwarp();
    } else {
        f5_epilogue(P, sh, tmem, Hs, warp, lane, (warp - F5_EPI_WARP0) >> 2);
    }
    tc5::fence_before_sync();
    __syncthreads();
    if (warp == F5_MMA_WARP) tc5::tmem_dealloc(tmem, tmem_cols);
}

// ------------------------------------------------------------------------------------------------ launcher (ffmlp.cu)
int f5_ffmlp_forward(const void *inputs, const void *weights, int64_t B, int input_dim, int num_layers, int activation,
                     void *forward_buffer, void *outputs, cudaStream_t stream) {
    const size_t w_bytes = (size_t)F5_HID * input_dim * 2 + (size_t)(num_layers - 1) * F5_HID * F5_HID * 2 + F5_OUT * F5_HID * 2;
    const size_t slot_bytes = (size_t)F5_ROWS * input_dim * 2 + F5_ROWS * F5_HID * 2;
    const size_t budget = 227 * 1024;
    if (1024 + w_bytes + slot_bytes > budget) return -1;   // the caller falls back to the one-tile-per-CTA kernel
    int slots = (int)((budget - 1024 - w_bytes) / slot_bytes);
    if (slots > F5_MAX_SLOTS) slots = F5_MAX_SLOTS;
    const size_t smem = 1024 + w_bytes + slots * slot_bytes;
    if (cudaFuncSetAttribute(k_ffmlp5_forward, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) {
        ngp_set_error("ngp_ffmlp_forward: cannot reserve %zu B of shared memory", smem);
        return NGP_ERR_CUDA;
    }
    uint32_t cols = 32;
    while (cols < (uint32_t)slots * F5_HID) cols <<= 1;
    int dev = 0, sms = NGP_SM_COUNT;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const int64_t tiles = ngp_div_up(B, F5_ROWS);
    F5Params P;
    P.x = (const __half *)inputs; P.w = (const __half *)weights; P.fwd_buf = (__half *)forward_buffer; P.out = (__half *)outputs;
    P.B = B; P.in_dim = input_dim; P.n_layers = num_layers; P.act = activation; P.slots = slots;
    k_ffmlp5_forward<<<(int)(tiles < sms ? tiles : sms), F5_THREADS, smem, stream>>>(P, cols);
    NGP_CHECK_LAUNCH("ngp_ffmlp_forward (warp-specialised)");
    return NGP_OK;
}

// ================================================================================================ fused encode -> FFMLP
// NeRFNetwork.density of the FFMLP variant (nerf/network_ff.py:76-89): h = sigma_net(encoder(x, bound)).  Unfused that is
// k_grid_forward -> [B, 32] fp32 features in HBM -> .half() -> FFMLP: 128 + 128 + 64 + 64 B per point of round trips between
// three launches.  Here the gather warps of the field kernels (cfield5.cuh: thread = point, 8 corners x 16 levels from the
// L2-resident fp32 table) write the features as fp16 straight into the FFMLP's chunk-major A-operand tile in shared memory
// and the roles above (MMA threads, epilogue warps) run the layers: 12 B in, 32 B out per point.  Features are computed in
// fp32 and rounded to fp16 once, exactly what `.half()` does to the encoder output in ffmlp.py:20-24.
// 4 tile slots (113 KB of shared memory with two hidden layers): the gathers are the bound and they need the L1 capacity the
// shared-memory carve-out takes away, so the kernel stays under 120 KB instead of filling the SM with slots.
#define FF5_GATHER_WARPS 8
#define FF5_MMA_WARP0 FF5_GATHER_WARPS
#define FF5_MMA_WARPS 4
#define FF5_EPI_WARP0 (FF5_MMA_WARP0 + FF5_MMA_WARPS)      // 12: a multiple of 4 (TMEM quadrant = warp % 4)
#define FF5_THREADS ((FF5_EPI_WARP0 + 4) * 32)             // 512
#define FF5_SLOTS 4

struct FF5Shared {
    F5Shared f;
    FqLevel lv[FQ_L];
    int all_safe;
};
static_assert(sizeof(FF5Shared) <= 1024, "FF5Shared must fit the first KB of the dynamic shared memory");

struct FF5Field {
    const float *x, *table;
    float bound;
    int align_corners;
};

template <bool SAFE>
__device__ __forceinline__ void ff5_gather(const F5Params &P, const FF5Field &G, FF5Shared *sh, unsigned char *Xs, int warp, int lane) {
    const int g = warp >> 2, q = warp & 3, row = 32 * q + lane;
    const int64_t ntiles = (P.B + F5_ROWS - 1) / F5_ROWS;
    for (int64_t j = 0;; ++j) {
        const int64_t n = 2 * j + g, tile = blockIdx.x + (int64_t)gridDim.x * n;   // group g gathers the CTA's tiles n = g, g + 2, ...
        if (tile >= ntiles) break;
        const int s = (int)(n & (FF5_SLOTS - 1));
        const int64_t r = n / FF5_SLOTS;
        const int64_t b = tile * F5_ROWS + row;
        float x01[3];
        const bool valid = fq_load_point(G.x, b, P.B, G.bound, x01);
        unsigned char *X = Xs + (size_t)s * F5_ROWS * 32 * 2;
#pragma unroll 1
        for (int lp = 0; lp < 8; ++lp) {
            float feat[4];
            c5_encode_pair<SAFE>(G.table, sh->lv, lp, x01, valid, G.align_corners, feat);
            if (lp == 0 && r > 0) tc5::mbar_wait_wd(tc5::smem_u32(&sh->f.bars[s].x_empty), (uint32_t)((r - 1) & 1));
            const __half2 a = __floats2half2_rn(feat[0], feat[1]), c = __floats2half2_rn(feat[2], feat[3]);
            uint2 u;
            u.x = *reinterpret_cast<const uint32_t *>(&a);
            u.y = *reinterpret_cast<const uint32_t *>(&c);
            // features 4 lp .. 4 lp + 3 of the row: 16-byte chunk lp / 2, its lower or upper half
            *reinterpret_cast<uint2 *>(X + tc5::chunk_off(F5_ROWS, row, lp >> 1) + (lp & 1) * 8) = u;
        }
        tc5::fence_async_smem();   // generic-proxy stores -> visible to the MMA's async-proxy reads
        tc5::mbar_arrive(tc5::smem_u32(&sh->f.bars[s].x_full));
    }
}

__global__ void __launch_bounds__(FF5_THREADS, 1)
k_ffield5_forward(F5Params P, FF5Field G, const int32_t *__restrict__ offsets, float S_, uint32_t H, uint32_t gridtype,
                  uint32_t tmem_cols) {
    extern __shared__ __align__(1024) unsigned char f5_smem[];
    FF5Shared *sh = reinterpret_cast<FF5Shared *>(f5_smem);
    const int n = P.n_layers;
    unsigned char *Win = f5_smem + 1024;
    unsigned char *Wh = Win + F5_HID * 32 * 2;
    unsigned char *Wout = Wh + (n - 1) * F5_HID * F5_HID * 2;
    unsigned char *Xs = Wout + F5_OUT * F5_HID * 2;
    unsigned char *Hs = Xs + (size_t)FF5_SLOTS * F5_ROWS * 32 * 2;
    const int tid = threadIdx.x, warp = __shfl_sync(0xffffffffu, tid >> 5, 0), lane = tid & 31;
    fq_stage_levels(sh->lv, &sh->all_safe, offsets, S_, H, gridtype, G.align_corners);
    if (tid == 0) {
        for (int s = 0; s < FF5_SLOTS; ++s) {
            tc5::mbar_init(tc5::smem_u32(&sh->f.bars[s].x_full), 128);
            tc5::mbar_init(tc5::smem_u32(&sh->f.bars[s].x_empty), 1);
            tc5::mbar_init(tc5::smem_u32(&sh->f.bars[s].d_full), 1);
            tc5::mbar_init(tc5::smem_u32(&sh->f.bars[s].a_ready), 128);
        }
        tc5::fence_mbar_init();
    }
    if (warp == FF5_MMA_WARP0) tc5::tmem_alloc(tc5::smem_u32(&sh->f.tmem), tmem_cols);
    f5_stage_weights(Win, P.w, F5_HID, 32);
    for (int l = 0; l < n - 1; ++l)
        f5_stage_weights(Wh + l * F5_HID * F5_HID * 2, P.w + F5_HID * 32 + l * F5_HID * F5_HID, F5_HID, F5_HID);
    f5_stage_weights(Wout, P.w + F5_HID * 32 + (n - 1) * F5_HID * F5_HID, F5_OUT, F5_HID);
    tc5::fence_async_smem();
    tc5::fence_before_sync();
    __syncthreads();
    tc5::fence_after_sync();
    const uint32_t tmem = __shfl_sync(0xffffffffu, sh->f.tmem, 0);
    if (warp < FF5_GATHER_WARPS) {
        if (sh->all_safe) ff5_gather<true>(P, G, sh, Xs, warp, lane);
        else ff5_gather<false>(P, G, sh, Xs, warp, lane);
    } else if (warp < FF5_EPI_WARP0) {
        if (lane == 0)
            f5_mma(P, &sh->f, tmem, tc5::smem_u32(Xs), tc5::smem_u32(Hs), tc5::smem_u32(Win), tc5::smem_u32(Wh), tc5::smem_u32(Wout),
                   warp - FF5_MMA_WARP0, FF5_MMA_WARPS);
        __syncwarp();
    } else {
        f5_epilogue(P, &sh->f, tmem, Hs, warp, lane, 0, 1);
    }
    tc5::fence_before_sync();
    __syncthreads();
    if (warp == FF5_MMA_WARP0) tc5::tmem_dealloc(tmem, tmem_cols);
}

extern "C" int ngp_field_query_ff_forward(const float *x, float bound, const float *embeddings, const int32_t *offsets, float S,
                                          uint32_t H, uint32_t gridtype, int align_corners, const void *weights, int num_layers,
                                          int activation, void *outputs, int64_t B, ngp_stream_t stream) {
    NGP_REQUIRE_DEVICE(x, "ngp_field_query_ff_forward");
    if (B == 0) return NGP_OK;
    NGP_REQUIRE(x && embeddings && offsets && weights && outputs, NGP_ERR_ARG, "ngp_field_query_ff_forward: null pointer");
    NGP_REQUIRE(bound > 0.0f && B >= 0, NGP_ERR_ARG, "ngp_field_query_ff_forward: bad bound / B");
    NGP_REQUIRE(((uintptr_t)embeddings & 15u) == 0 && ((uintptr_t)weights & 15u) == 0 && ((uintptr_t)outputs & 31u) == 0, NGP_ERR_ARG,
                "ngp_field_query_ff_forward: embeddings / weights must be 16-byte, outputs 32-byte aligned");
    NGP_REQUIRE(num_layers >= 2 && num_layers <= 6, NGP_ERR_UNSUPPORTED, "ngp_field_query_ff_forward: num_layers must be in 2..6 (got %d)",
                num_layers);
    NGP_REQUIRE(activation == 0 || activation == 6, NGP_ERR_UNSUPPORTED,
                "ngp_field_query_ff_forward: hidden activation must be relu (0) or none (6); got %d", activation);
    const size_t w_bytes = (size_t)F5_HID * 32 * 2 + (size_t)(num_layers - 1) * F5_HID * F5_HID * 2 + F5_OUT * F5_HID * 2;
    const size_t smem = 1024 + w_bytes + (size_t)FF5_SLOTS * (F5_ROWS * 32 * 2 + F5_ROWS * F5_HID * 2);
    if (cudaFuncSetAttribute(k_ffield5_forward, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) {
        ngp_set_error("ngp_field_query_ff_forward: cannot reserve %zu B of shared memory", smem);
        return NGP_ERR_CUDA;
    }
    int dev = 0, sms = NGP_SM_COUNT;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const int64_t tiles = ngp_div_up(B, F5_ROWS);
    F5Params P;
    P.x = nullptr; P.w = (const __half *)weights; P.fwd_buf = nullptr; P.out = (__half *)outputs;
    P.B = B; P.in_dim = 32; P.n_layers = num_layers; P.act = activation; P.slots = FF5_SLOTS;
    FF5Field G;
    G.x = x; G.table = embeddings; G.bound = bound; G.align_corners = align_corners;
    // 512 TMEM columns although 256 are used: a CTA that owns them all keeps a second CTA off the SM (one persistent CTA
    // per SM is the design, and the shared-memory request alone would admit two)
    k_ffield5_forward<<<(int)(tiles < sms ? tiles : sms), FF5_THREADS, smem, (cudaStream_t)stream>>>(P, G, offsets, S, H, gridtype, 512u);
    NGP_CHECK_LAUNCH("ngp_field_query_ff_forward");
    return NGP_OK;
}
